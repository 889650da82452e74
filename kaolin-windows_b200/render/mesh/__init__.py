# mirrors kaolin/render/mesh/__init__.py:1-4 for the hot path only
from .rasterization import *
from .dibr import *
from .utils import *
from .deftet import *
from . import rasterization, dibr, utils, deftet
