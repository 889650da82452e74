# mirrors kaolin/ops/__init__.py for the one module next to the DIB-R path
from . import mesh
