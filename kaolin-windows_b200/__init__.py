"""kaolin-windows_b200 -- B200-native (sm_100a) drop-in for Kaolin's DIB-R differentiable
rasterization hot path: ``render.mesh.rasterize(backend='cuda')``,
``render.mesh.dibr_soft_mask`` and ``render.mesh.dibr_rasterization``, forward and
backward, over hand-written CUDA kernels behind a C ABI (include/dibr_b200.h).

Next to the path: ``render.mesh.prepare_vertices`` and ``ops.mesh.index_vertices_by_faces``
(the step that feeds it), ``render.mesh.texture_mapping`` and ``metrics.render.mask_iou``
(the step that consumes it), and the sibling op ``render.mesh.deftet_sparse_render``.
Nothing else of Kaolin is provided.  The directory name is not a Python identifier; import
it as ``kaolin_windows_b200`` (a small alias package at the repository root) or with
``importlib.import_module("kaolin-windows_b200")``.
"""
from . import _lib
from . import _C
from . import render
from . import ops
from . import metrics
from . import parallel
from ._lib import build

__version__ = "0.1.0"
__all__ = ["render", "ops", "metrics", "parallel", "_C", "build"]
