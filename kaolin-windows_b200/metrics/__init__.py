# mirrors kaolin/metrics/__init__.py for the one module next to the DIB-R path
from . import render
