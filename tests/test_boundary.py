"""CPU-side checks of the drop-in boundary: the C-ABI library loads and exports every
symbol include/dibr_b200.h declares, the planning entry points answer without a GPU, and
the Python mirror of the reference API keeps its signatures, defaults and error behaviour.
No compute call is made here."""
import inspect
import os
import re

import pytest
import torch

import kaolin_windows_b200 as kal
from kaolin_windows_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "dibr_b200.h")).read()
    return re.findall(r"^DIBR_API\s+[\w\s\*]+?\b(dibr_\w+)\s*\(", src, flags=re.M)


def test_library_exports_every_declared_symbol():
    names = _declared_symbols()
    assert len(names) == 38 and "dibr_deftet_fwd" in names and "dibr_render_fwd" in names and "dibr_face_setup_fwd" in names
    h = _lib.lib()
    for n in names:
        assert hasattr(h, n), n
    assert set(names) == set(_lib.SIGNATURES)
    assert h.dibr_version() >= 100


def test_header_argument_counts_match_binding():
    src = open(os.path.join(ROOT, "include", "dibr_b200.h")).read()
    for name, (_, args) in _lib.SIGNATURES.items():
        m = re.search(r"^DIBR_API[^;(]*\b%s\s*\(([^;]*?)\)\s*;" % name, src, flags=re.S | re.M)
        assert m, name
        params = m.group(1).strip()
        n = 0 if params in ("void", "") else params.count(",") + 1
        assert n == len(args), (name, n, len(args))


def test_workspace_planning_without_gpu():
    h = _lib.lib()
    fixed = h.dibr_workspace_fixed_bytes(8, 70000, 512, 512)
    assert fixed > 0 and fixed % 256 == 0
    # raster entries are 48-byte face records (4-byte ids with DIBR_B200_RASTER_REC=0), soft
    # entries 4-byte face ids; each list area is 256-byte aligned
    rb = 4 if os.environ.get("DIBR_B200_RASTER_REC", "1")[:1] == "0" else 48
    up = lambda v: (v + 255) // 256 * 256
    assert h.dibr_workspace_bytes_for(8, 70000, 512, 512, 10, 20, 0, 0) == fixed + up(rb * 10) + up(4 * 20)
    assert h.dibr_workspace_fixed_bytes(1, 10, 0, 16) == _lib.DIBR_E_INVALID
    assert "invalid sizes" in _lib.last_error()
    # larger images use larger bins, never more than 1024 per view
    assert h.dibr_workspace_fixed_bytes(16, 1000000, 2048, 2048) < (1 << 30)


def test_public_signatures_match_reference():
    """Names, order and defaults of kaolin/render/mesh/rasterization.py:390-398 and
    kaolin/render/mesh/dibr.py:75-76,119-122."""
    m = kal.render.mesh
    sig = inspect.signature(m.rasterize)
    assert list(sig.parameters) == ["height", "width", "face_vertices_z", "face_vertices_image",
                                    "face_features", "valid_faces", "multiplier", "eps", "backend"]
    assert sig.parameters["backend"].default == "cuda"
    assert sig.parameters["multiplier"].default is None and sig.parameters["eps"].default is None
    sig = inspect.signature(m.dibr_soft_mask)
    assert [(k, v.default) for k, v in sig.parameters.items()][2:] == [
        ("sigmainv", 7000), ("boxlen", 0.02), ("knum", 30), ("multiplier", 1000.)]
    sig = inspect.signature(m.dibr_rasterization)
    assert list(sig.parameters) == ["height", "width", "face_vertices_z", "face_vertices_image",
                                    "face_features", "face_normals_z", "sigmainv", "boxlen", "knum",
                                    "multiplier", "eps", "rast_backend",
                                    "cuda_graph"]      # extension, keyword, default off
    assert sig.parameters["cuda_graph"].default is False
    for n in ("packed_rasterize_forward_cuda", "rasterize_backward_cuda",
              "dibr_soft_mask_forward_cuda", "dibr_soft_mask_backward_cuda"):
        assert callable(getattr(kal._C.render.mesh, n))      # bindings.cpp:76-81


def test_no_cpu_fallback_and_error_behaviour():
    z = torch.zeros(1, 2, 3)
    img = torch.zeros(1, 2, 3, 2)
    feat = torch.zeros(1, 2, 3, 1)
    with pytest.raises(RuntimeError, match="CUDA"):
        kal.render.mesh.rasterize(8, 8, z, img, feat)
    with pytest.raises(RuntimeError, match="CUDA"):
        kal.render.mesh.dibr_soft_mask(img, torch.zeros(1, 8, 8, dtype=torch.long))
    with pytest.raises(ValueError, match="not a valid backend"):     # rasterization.py:495-497
        kal.render.mesh.rasterize(8, 8, z, img, feat, backend="opengl")
    with pytest.raises(ValueError, match="nvdiffrast"):
        kal.render.mesh.rasterize(8, 8, z, img, feat, backend="nvdiffrast")
    with pytest.raises(RuntimeError):
        kal._C.render.mesh.packed_rasterize_forward_cuda(
            8, 8, z[0], img[0], torch.zeros(2, 4), feat[0], torch.zeros(2, dtype=torch.long), 1000, 1e-8)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "kaolin-windows_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".sh")):
                assert "oracle" not in open(os.path.join(dp, f)).read(), os.path.join(dp, f)


def test_shard_range_partitions_views():
    from kaolin_windows_b200.parallel import shard_range
    for n, ws in [(64, 8), (64, 2), (256, 8), (7, 4), (3, 8)]:
        spans = [shard_range(n, r, ws) for r in range(ws)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
        assert max(hi - lo for lo, hi in spans) - min(hi - lo for lo, hi in spans) <= 1


def test_deftet_signature_and_planning_without_gpu():
    """kaolin/render/mesh/deftet.py:342-343: names, order, defaults; the workspace bound and
    the argument checks answer without a GPU; CPU tensors are refused (no CPU path)."""
    sig = inspect.signature(kal.render.mesh.deftet_sparse_render)
    assert list(sig.parameters) == ["pixel_coords", "render_ranges", "face_vertices_z",
                                    "face_vertices_image", "face_features", "knum", "eps"]
    assert sig.parameters["knum"].default == 300 and sig.parameters["eps"].default == 1e-8
    h = _lib.lib()
    n = h.dibr_deftet_workspace_bytes(8, 70000, 16384)
    assert n > 0 and n % 256 == 0
    # bound: 16 list entries + 1 wide-list entry per face and view, two ints per cell
    assert n < 8 * 70000 * 4 * 17 + 8 * 2 * 4 * (4 * 16384 + 16) + (1 << 16)
    assert h.dibr_deftet_workspace_bytes(-1, 1, 1) == _lib.DIBR_E_INVALID
    assert h.dibr_deftet_fwd(7, 1, 1, 1, 1, 1, None, None, None, None, None, 1e-8, None, None,
                             None, None, 0, None) == _lib.DIBR_E_INVALID
    assert "dtype" in _lib.last_error()
    z = torch.zeros(1, 2, 3)
    with pytest.raises(RuntimeError):
        kal.render.mesh.deftet_sparse_render(torch.zeros(1, 4, 2), torch.zeros(1, 4, 2), z,
                                             torch.zeros(1, 2, 3, 2), torch.zeros(1, 2, 3, 1), 3)


def test_peer_exchange_argument_checks_without_gpu():
    import ctypes
    h = _lib.lib()
    pads = (ctypes.c_int64 * 2)(0, 0)
    assert h.dibr_peer_signal(0, 0, pads, 0, None) == _lib.DIBR_E_INVALID       # world < 1
    assert h.dibr_peer_signal(0, 2, pads, 0, None) == _lib.DIBR_E_INVALID       # null pads
    assert "signal pad" in _lib.last_error()
    assert h.dibr_peer_allreduce(3, 2, pads, pads, 8, None, None, None) == _lib.DIBR_E_INVALID
