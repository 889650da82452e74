"""float64 on the B200: the reference dispatches float and double
(rasterization_cuda.cu:30-39) and parametrises its tests over both
(test_rasterization.py:32, test_dibr.py:37).  The double kernels (csrc/dibr_f64.cu, reached
through the public API and the `_C.render.mesh` shim) against the reference's own kernels
compiled into oracle/_ref and run in double on the same inputs: face indices and the
per-pixel soft-mask state bit-exact, values and gradients to 1e-10 relative."""
import pytest
import torch

import workloads as wl
from oracle import dibr_oracle as orc

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module")
def kal():
    import kaolin_windows_b200 as k
    return k


@pytest.fixture(scope="module")
def ref():
    return orc.load_ref_ops()


def rel(a, b):
    a, b = a.detach(), b.detach()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-300))


def _geo(batch_size, flip):
    g = wl.fixture_geometry(batch_size=batch_size, flip=flip, dtype=torch.float64)
    return {k: v.to(DEV) for k, v in g.items()}


@pytest.mark.parametrize("batch_size", [1, 3])
@pytest.mark.parametrize("flip", [False, True])
@pytest.mark.parametrize("with_valid_faces", [False, True])
@pytest.mark.parametrize("hw", [(35, 31), (128, 96)])
def test_rasterize_double(kal, ref, batch_size, flip, with_valid_faces, hw):
    H, W = hw
    g = _geo(batch_size, flip)
    z, img, uvs = g["face_vertices_z"], g["face_vertices_image"], g["face_uvs"]
    feat = torch.cat([uvs, torch.ones_like(uvs[..., 1:])], dim=-1).contiguous()
    valid = None
    if with_valid_faces:
        zz = z.reshape(batch_size, -1)
        mid = ((zz.min(1, keepdim=True)[0] + zz.max(1, keepdim=True)[0]) / 2.).unsqueeze(-1)
        valid = torch.all(z < mid, dim=-1)
    a = img.clone().requires_grad_(True)
    b = feat.clone().requires_grad_(True)
    out, idx = kal.render.mesh.rasterize(H, W, z, a, b, valid)
    assert out.dtype == torch.float64 and idx.dtype == torch.long
    want_out, want_idx, ctx = orc.rasterize_forward(H, W, z, img, feat, valid, ops=ref)
    assert torch.equal(idx, want_idx)
    assert rel(out, want_out) <= 1e-12
    cot = torch.rand(out.shape, dtype=torch.float64, device=DEV,
                     generator=torch.Generator(device=DEV).manual_seed(3))
    out.backward(cot)
    gi, gf = orc.rasterize_backward_from_ctx(ctx, cot, ops=ref)
    assert rel(a.grad, gi) <= 1e-10 and rel(b.grad, gf) <= 1e-10


@pytest.mark.parametrize("sigmainv,boxlen,knum", [(7000., 0.02, 30), (70., 0.2, 5), (7000., 0.02, 1)])
@pytest.mark.parametrize("hw", [(35, 31), (96, 128)])
def test_dibr_soft_mask_and_rasterization_double(kal, ref, sigmainv, boxlen, knum, hw):
    H, W = hw
    g = _geo(2, False)
    z, img, uvs = g["face_vertices_z"], g["face_vertices_image"], g["face_uvs"]
    feat = torch.cat([uvs, torch.ones_like(uvs[..., 1:])], dim=-1).contiguous()
    nz = g["face_normals_z"]
    a = img.clone().requires_grad_(True)
    b = feat.clone().requires_grad_(True)
    out, mask, idx = kal.render.mesh.dibr_rasterization(H, W, z, a, b, nz, sigmainv, boxlen, knum)
    assert out.dtype == torch.float64 and mask.dtype == torch.float64
    w_out, w_mask, w_idx, ctxs = orc.dibr_rasterization_forward(H, W, z, img, feat, nz, sigmainv,
                                                              boxlen, knum, ops=ref)
    assert torch.equal(idx, w_idx)
    assert rel(out, w_out) <= 1e-12 and float((mask - w_mask).abs().max()) <= 1e-12
    gen = torch.Generator(device=DEV).manual_seed(4)
    g1 = torch.rand(out.shape, dtype=torch.float64, device=DEV, generator=gen)
    g2 = torch.rand(mask.shape, dtype=torch.float64, device=DEV, generator=gen)
    torch.autograd.backward((out, mask), (g1, g2))
    gi, gf = orc.dibr_rasterization_backward(ctxs, g1, g2, ops=ref)
    assert rel(a.grad, gi) <= 1e-10 and rel(b.grad, gf) <= 1e-10
    # the raw op's dense state, as the reference's own test reads it (test_dibr.py:123-130)
    scaled = (img * 1000.).contiguous()
    lo = scaled.min(dim=-2)[0] - boxlen * 1000.
    hi = scaled.max(dim=-2)[0] + boxlen * 1000.
    boxes = torch.cat([lo, hi], dim=-1).contiguous()
    got = kal._C.render.mesh.dibr_soft_mask_forward_cuda(scaled, boxes, idx, sigmainv, knum, 1000.)
    want = ref.dibr_soft_mask_forward_cuda(scaled, boxes, idx, sigmainv, knum, 1000.)
    assert torch.equal(got[2], want[2]) and torch.equal(got[3], want[3])
    assert float((got[0] - want[0]).abs().max()) <= 1e-12
    assert float((got[1] - want[1]).abs().max()) <= 1e-12


def test_mixed_dtypes_raise(kal):
    g = _geo(1, False)
    z, img, uvs = g["face_vertices_z"], g["face_vertices_image"], g["face_uvs"]
    with pytest.raises(RuntimeError):
        kal.render.mesh.rasterize(16, 16, z, img.float(), uvs)


def test_config1_double(kal, ref):
    """BASELINE config 1 (pizza.obj, one 256 x 256 view) in double, forward and backward."""
    inp = wl.make_config("cfg1_pizza_256", device=DEV)
    H, W = inp["H"], inp["W"]
    z, img, feat, nz = (inp[k].double().contiguous() for k in
                        ("face_vertices_z", "face_vertices_image", "face_features",
                         "face_normals_z"))
    a = img.clone().requires_grad_(True)
    b = feat.clone().requires_grad_(True)
    out, mask, idx = kal.render.mesh.dibr_rasterization(H, W, z, a, b, nz)
    w_out, w_mask, w_idx, ctxs = orc.dibr_rasterization_forward(H, W, z, img, feat, nz, ops=ref)
    assert torch.equal(idx, w_idx) and int((idx >= 0).sum()) > 1000
    assert rel(out, w_out) <= 1e-12 and float((mask - w_mask).abs().max()) <= 1e-12
    g1, g2 = torch.ones_like(out), torch.ones_like(mask)
    torch.autograd.backward((out, mask), (g1, g2))
    gi, gf = orc.dibr_rasterization_backward(ctxs, g1, g2, ops=ref)
    assert rel(a.grad, gi) <= 1e-10 and rel(b.grad, gf) <= 1e-10
