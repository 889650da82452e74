"""Importable alias of the ``kaolin-windows_b200/`` package directory (whose name is not a
Python identifier): ``import kaolin_windows_b200 as kal`` then
``kal.render.mesh.dibr_rasterization(...)``."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                    "kaolin-windows_b200")
_spec = importlib.util.spec_from_file_location(
    __name__, os.path.join(_dir, "__init__.py"), submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules[__name__] = _mod
_spec.loader.exec_module(_mod)
