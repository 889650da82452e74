"""dibr_rasterize_fwd_static: the allocation-free variant of the rasterizer forward (scratch
workspace in a file-scope device array of DIBR_STATIC_SCRATCH_BYTES inside the library; no
counterpart in the reference) against the
existing entry point dibr_rasterize_fwd with a caller-allocated workspace.

Both are called through the C ABI (ctypes).  The forward is bit-exact in this repository
(selected_face_idx, interpolated features and weights compared with torch.equal everywhere),
so that is the tolerance here too.  Two calls in a row on different inputs and image sizes:
the array keeps the first call's contents and the second must not see them.  The file name
makes pytest collect it last."""
import ctypes

import pytest
import torch

import workloads as wl

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
STATIC_BYTES = 16 * 1024 * 1024      # DIBR_STATIC_SCRATCH_BYTES in include/dibr_b200.h


def _p(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _outputs(B, H, W, D):
    # poisoned, so that an element the call fails to write shows up
    return (torch.full((B, H, W), -7, dtype=torch.long, device=DEV),
            torch.full((B, H, W, D), float("nan"), device=DEV),
            torch.full((B, H, W, 3), float("nan"), device=DEV))


def _existing(lib, _lib, H, W, z, img, feat, nz):
    B, F = z.shape[:2]
    D = feat.shape[-1]
    idx, out, wts = _outputs(B, H, W, D)
    cap = 4 * B * F + 4096
    n = int(lib.dibr_workspace_bytes_for(B, F, H, W, cap, 0, 0, 0))
    ws = torch.empty(n, dtype=torch.uint8, device=DEV)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.dibr_rasterize_fwd(B, F, H, W, D, _p(z), _p(img), _p(feat), None, _p(nz), 0, None,
                                None, 1000.0, 1e-8, _p(idx), _p(out), _p(wts), _p(ws), n, cap,
                                None, st)
    _lib.check(rc, "dibr_rasterize_fwd")
    torch.cuda.synchronize()
    return idx, out, wts


def _static(lib, _lib, H, W, z, img, feat, nz):
    B, F = z.shape[:2]
    D = feat.shape[-1]
    idx, out, wts = _outputs(B, H, W, D)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.dibr_rasterize_fwd_static(B, F, H, W, D, _p(z), _p(img), _p(feat), None, _p(nz),
                                       1000.0, 1e-8, _p(idx), _p(out), _p(wts), None, st)
    _lib.check(rc, "dibr_rasterize_fwd_static")
    torch.cuda.synchronize()
    return idx, out, wts


def test_static_scratch_variant_matches_existing_entry_point_twice():
    from kaolin_windows_b200 import _lib
    lib = _lib.lib()
    torch.cuda.set_device(0)
    g = wl.fixture_geometry(batch_size=3)
    a = tuple(g[k].to(DEV).contiguous() for k in ("face_vertices_z", "face_vertices_image",
                                                  "face_uvs", "face_normals_z"))
    c = wl.make_config("cfg1_pizza_256", device=DEV)
    b = tuple(c[k].contiguous() for k in ("face_vertices_z", "face_vertices_image",
                                          "face_features", "face_normals_z"))
    # two calls in a row (then the first again): different meshes, batch and image sizes
    for (H, W, inp) in ((96, 80, a), (256, 256, b), (96, 80, a)):
        want = _existing(lib, _lib, H, W, *inp)
        got = _static(lib, _lib, H, W, *inp)
        assert int((want[0] >= 0).sum()) > 500
        for x, y in zip(got, want):
            assert x.dtype == y.dtype and torch.equal(x, y)


def test_static_scratch_variant_rejects_sizes_beyond_the_array():
    from kaolin_windows_b200 import _lib
    lib = _lib.lib()
    torch.cuda.set_device(0)
    # BASELINE config 2 (8 x 69 938 faces at 512 x 512) needs more than 16 MiB of tables and
    # lists; rejected before anything is looked at or launched (null pointers on purpose)
    assert int(lib.dibr_workspace_fixed_bytes(8, 69938, 512, 512)) + 48 * 2 * 8 * 69938 > STATIC_BYTES
    rc = lib.dibr_rasterize_fwd_static(8, 69938, 512, 512, 3, None, None, None, None, None,
                                       1000.0, 1e-8, None, None, None, None, None)
    assert rc == _lib.DIBR_E_INVALID
    msg = _lib.last_error()
    assert "DIBR_STATIC_SCRATCH_BYTES" in msg and str(STATIC_BYTES) in msg
