"""GPU parity tests: the sm_100a kernels (through the C ABI) against

* the reference's own CUDA kernels compiled unmodified into oracle/_ref (authoritative),
* the reference's golden tensors / naive torch rasterizer (tests/golden/*.npz),
* the CPU oracle (fp64 instantiation = low-noise gradients).

Bars (BASELINE.json): selected_face_idx bit-exact; features and soft mask <= 1e-5;
gradients <= 1e-4 relative (to the largest reference gradient magnitude of the tensor).
The parametrisation mirrors tests/python/kaolin/render/mesh/test_rasterization.py and
test_dibr.py of the reference.
"""
import os

import numpy as np
import pytest
import torch

import parity_util as pu
import workloads as wl
from oracle import dibr_oracle as orc

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
H, W = 35, 31
GOLDEN = wl.GOLDEN


@pytest.fixture(scope="module")
def kal():
    import kaolin_windows_b200 as k
    return k


@pytest.fixture(scope="module")
def ref():
    return orc.load_ref_ops()


def _npz(name):
    return np.load(os.path.join(GOLDEN, name))


def _g(npz, name, dtype=None):
    t = torch.from_numpy(np.asarray(npz[name])).to(DEV)
    return t.to(dtype) if dtype is not None else t


def rel_err(a, b):
    """max |a-b| relative to the largest magnitude of the reference tensor."""
    a, b = a.double(), b.double()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def grad_ok(mine, ref32, ref64, what="grad_face_vertices_image"):
    """Per-face gradient bar of tests/parity_util.py: every face within 1e-4 of the reference
    run in fp64 (relative to the face's own gradient, floored at 1e-3 of the tensor's
    largest), or -- where the reference's own fp32 arithmetic is itself further than that
    from fp64 (it forms 1 - soft_mask and 1 - prob in float) -- no further from fp64 than
    1.5x the reference's fp32 result, capped at 1e-3."""
    st = pu.face_grad_check(what, mine, ref32.to(mine.device), ref64.to(mine.device))
    if not st["ok"]:
        print("PARITY-FAIL", st)
    return st["ok"]


def geo(batch_size, flip):
    g = wl.fixture_geometry(batch_size=batch_size, flip=flip)
    return {k: v.to(DEV) for k, v in g.items()}


# ---------------------------------------------------------------------- rasterize
@pytest.mark.parametrize("batch_size", [1, 3])
@pytest.mark.parametrize("flip", [False, True])
@pytest.mark.parametrize("with_valid_faces", [False, True])
class TestRasterize:
    def _inputs(self, batch_size, flip, with_valid_faces):
        g = geo(batch_size, flip)
        naive = _npz("rasterize_naive.npz")
        valid = None
        if with_valid_faces:
            valid = _g(naive, "float32_flip%d_validmask" % flip)[:batch_size]
        uvs = g["face_uvs"]
        feats = [uvs, torch.ones_like(uvs[..., 1:])]
        return g["face_vertices_z"], g["face_vertices_image"], feats, valid, \
            "float32_flip%d_valid%d" % (flip, with_valid_faces), naive

    def test_forward(self, kal, ref, batch_size, flip, with_valid_faces):
        z, img, feats, valid, key, naive = self._inputs(batch_size, flip, with_valid_faces)
        (uv, ones), idx = kal.render.mesh.rasterize(H, W, z, img, feats, valid_faces=valid)
        out = torch.cat([uv, ones], dim=-1)
        # reference's own kernels
        r_out, r_idx, _ = orc.rasterize_forward(H, W, z, img, feats, valid, ops=ref)
        assert torch.equal(idx, r_idx)
        assert torch.equal(out, r_out)          # same arithmetic -> same bits
        # reference's naive torch rasterizer (test_rasterization.py:148-158)
        assert torch.equal(idx, _g(naive, key + "_idx").long()[:batch_size])
        assert torch.allclose(out, _g(naive, key + "_feat")[:batch_size], rtol=1e-5, atol=1e-5)
        assert idx.dtype == torch.long and (idx >= 0).sum() > 100

    def test_backward(self, kal, ref, batch_size, flip, with_valid_faces):
        z, img, feats, valid, key, naive = self._inputs(batch_size, flip, with_valid_faces)
        z = z.clone().requires_grad_(True)
        img = img.clone().requires_grad_(True)
        f0 = feats[0].clone().requires_grad_(True)
        f1 = feats[1].clone().requires_grad_(True)
        (uv, ones), idx = kal.render.mesh.rasterize(H, W, z, img, [f0, f1], valid_faces=valid)
        out = torch.cat([uv, ones], dim=-1)
        grad_out = _g(naive, key + "_grad_out", torch.float32)[:batch_size].contiguous()
        out.backward(grad_out)
        assert z.grad is None                    # rasterization.py:387-388
        gf = torch.cat([f0.grad, f1.grad], dim=-1)
        # tolerances of test_rasterization.py:232-236 against the naive oracle's autograd
        assert torch.allclose(img.grad, _g(naive, key + "_grad_img", torch.float32)[:batch_size],
                              rtol=1e-3, atol=1e-2)
        assert torch.allclose(gf, _g(naive, key + "_grad_feat", torch.float32)[:batch_size],
                              rtol=1e-3, atol=1e-3)
        # per-face 1e-4 against the CPU oracle of the reference kernel run in fp64 on the
        # state its fp32 forward saved (same selected faces); its fp32 run beside it
        _, _, ctx32 = orc.rasterize_forward(H, W, z.detach().cpu(), img.detach().cpu(),
                                            [f0.detach().cpu(), f1.detach().cpu()],
                                            None if valid is None else valid.cpu())
        gi32, gf32 = orc.rasterize_backward_from_ctx(ctx32, grad_out.cpu())
        ctx64 = orc.ctxs_to_double((ctx32,))[0]
        gi64, gf64 = orc.rasterize_backward_from_ctx(ctx64, grad_out.double().cpu())
        assert grad_ok(img.grad.cpu(), gi32, gi64)
        assert grad_ok(gf.cpu(), gf32, gf64, "grad_face_features")


# ---------------------------------------------------------------- soft mask, goldens
def _large_bboxes(img_scaled, boxlen, multiplier):
    pmin = torch.min(img_scaled, dim=-2)[0]
    pmax = torch.max(img_scaled, dim=-2)[0]
    return torch.cat([pmin - boxlen * multiplier, pmax + boxlen * multiplier], dim=-1)


@pytest.mark.parametrize("sigmainv", [7000, 70])
@pytest.mark.parametrize("boxlen", [0.02, 0.2])
class TestSimpleDibrSoftMask:
    @pytest.fixture
    def inputs(self, kal):
        img, z = wl.simple_soft_mask_fixture()
        img, z = img.to(DEV), z.to(DEV)
        feat = torch.zeros(z.shape + (1,), device=DEV)
        _, face_idx = kal.render.mesh.rasterize(H, W, z, img, feat)
        return img, face_idx

    @pytest.mark.parametrize("multiplier", [1000, 100, 1])
    @pytest.mark.parametrize("knum", [30, 20])
    def test_C_forward(self, kal, inputs, sigmainv, boxlen, knum, multiplier):
        img, face_idx = inputs
        gt = _npz("dibr_simple.npz")
        key = "%d_%d_%d_%s" % (H, W, sigmainv, boxlen)
        img_s = img * multiplier
        mask, prob, cidx, typ = kal._C.render.mesh.dibr_soft_mask_forward_cuda(
            img_s, _large_bboxes(img_s, boxlen, multiplier), face_idx, sigmainv, knum, multiplier)
        assert torch.allclose(mask, _g(gt, "soft_mask_" + key, torch.float32), atol=1e-5, rtol=1e-5)
        assert torch.equal(cidx, _g(gt, "close_face_idx_" + key).long()[..., :knum])
        assert torch.allclose(prob, _g(gt, "close_face_dist_" + key, torch.float32)[..., :knum],
                              atol=1e-5, rtol=1e-5)
        assert torch.equal(typ, _g(gt, "close_face_dist_type_" + key)[..., :knum])

    @pytest.mark.parametrize("multiplier", [1000, 100, 1])
    @pytest.mark.parametrize("knum", [30, 20])
    def test_forward_backward(self, kal, inputs, sigmainv, boxlen, knum, multiplier):
        img, face_idx = inputs
        gt = _npz("dibr_simple.npz")
        key = "%d_%d_%d_%s" % (H, W, sigmainv, boxlen)
        img = img.clone().requires_grad_(True)
        mask = kal.render.mesh.dibr_soft_mask(img, face_idx, sigmainv, boxlen, knum, multiplier)
        assert torch.allclose(mask, _g(gt, "soft_mask_" + key, torch.float32), atol=1e-5, rtol=1e-5)
        shifted = torch.nn.functional.pad(face_idx != -1, (0, 5))[..., 5:]
        orc.mask_iou(mask, shifted).backward()
        assert torch.allclose(img.grad, _g(gt, "grad_face_vertices_image_" + key, torch.float32),
                              rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("batch_size", [1, 3])
@pytest.mark.parametrize("flip", [False, True])
@pytest.mark.parametrize("sigmainv", [7000, 70])
@pytest.mark.parametrize("boxlen", [0.02, 0.01])
class TestDibrSoftMask:
    @pytest.fixture
    def inputs(self, kal, batch_size, flip):
        g = geo(batch_size, flip)
        z, img = g["face_vertices_z"], g["face_vertices_image"]
        feat = torch.zeros(z.shape + (1,), device=DEV)
        _, face_idx = kal.render.mesh.rasterize(H, W, z, img, feat)
        return img, face_idx

    @pytest.mark.parametrize("multiplier", [1000, 100])
    @pytest.mark.parametrize("knum", [30, 40])
    def test_C_forward(self, kal, ref, inputs, batch_size, sigmainv, boxlen, knum, multiplier):
        img, face_idx = inputs
        gt = _npz("dibr_sphere.npz")
        key = "%d_%d_%d_%s" % (H, W, sigmainv, boxlen)
        b = batch_size
        img_s = img * multiplier
        bbox = _large_bboxes(img_s, boxlen, multiplier)
        mask, prob, cidx, typ = kal._C.render.mesh.dibr_soft_mask_forward_cuda(
            img_s, bbox, face_idx, sigmainv, knum, multiplier)
        assert torch.allclose(mask, _g(gt, "soft_mask_" + key, torch.float32)[:b],
                              atol=1e-5, rtol=1e-5)
        assert torch.equal(cidx, _g(gt, "close_face_idx_" + key).long()[:b, ..., :knum])
        assert torch.allclose(prob, _g(gt, "close_face_dist_" + key, torch.float32)[:b, ..., :knum],
                              atol=1e-5, rtol=1e-5)
        diff = typ != _g(gt, "close_face_dist_type_" + key)[:b, ..., :knum]
        assert torch.sum(diff) / diff.numel() <= 0.01           # test_dibr.py:340-341
        # and exactly what the reference's kernel computes on this GPU
        r_mask, r_prob, r_cidx, r_typ = ref.dibr_soft_mask_forward_cuda(
            img_s, bbox, face_idx, sigmainv, knum, multiplier)
        assert torch.equal(cidx, r_cidx) and torch.equal(typ, r_typ)
        assert torch.allclose(prob, r_prob, rtol=1e-6, atol=1e-7)
        assert torch.allclose(mask, r_mask, rtol=1e-6, atol=1e-7)

    @pytest.mark.parametrize("multiplier", [1000, 100, 1])
    @pytest.mark.parametrize("knum", [30, 40])
    def test_forward_backward(self, kal, ref, inputs, batch_size, sigmainv, boxlen, knum,
                              multiplier):
        img, face_idx = inputs
        gt = _npz("dibr_sphere.npz")
        key = "%d_%d_%d_%s" % (H, W, sigmainv, boxlen)
        img = img.clone().requires_grad_(True)
        mask = kal.render.mesh.dibr_soft_mask(img, face_idx, sigmainv, boxlen, knum, multiplier)
        if multiplier >= 100:   # the reference checks the mask golden for 1000 and 100 only
            assert torch.allclose(mask, _g(gt, "soft_mask_" + key, torch.float32)[:batch_size],
                                  atol=1e-5, rtol=1e-5)
        shifted = torch.nn.functional.pad(face_idx != -1, (0, 5))[..., 5:]
        orc.mask_iou(mask, shifted).backward()
        gtg = _g(gt, "grad_face_vertices_image_" + key, torch.float32)[:batch_size]
        assert torch.allclose(img.grad, gtg, rtol=1e-1, atol=1e-1)  # test_dibr.py:390-394
        # tight: against the reference kernels in fp32 (same arithmetic as ours up to the
        # order of the atomics) and in fp64 on the state the fp32 forward saved
        m32, ctx32 = orc.dibr_soft_mask_forward_py(img.detach(), face_idx, sigmainv, boxlen,
                                                   knum, multiplier, ops=ref)
        assert torch.allclose(mask, m32, rtol=1e-6, atol=1e-7)
        m32 = m32.clone().requires_grad_(True)
        orc.mask_iou(m32, shifted).backward()
        g32 = orc.dibr_soft_mask_backward_from_ctx(ctx32, m32.grad, ops=ref)
        ctx64 = orc.ctxs_to_double((ctx32,))[0]
        m64 = ctx64["soft_mask"].clone().requires_grad_(True)
        orc.mask_iou(m64, shifted).backward()
        g64 = orc.dibr_soft_mask_backward_from_ctx(ctx64, m64.grad, ops=ref)
        assert grad_ok(img.grad, g32, g64), (rel_err(img.grad, g64), rel_err(img.grad, g32),
                                             rel_err(g32, g64))
        # and the dense-state shim backward
        img_s = img.detach() * multiplier
        mk, prob, cidx, typ = kal._C.render.mesh.dibr_soft_mask_forward_cuda(
            img_s, _large_bboxes(img_s, boxlen, multiplier), face_idx, sigmainv, knum, multiplier)
        gm = mask.detach().clone().requires_grad_(True)
        orc.mask_iou(gm, shifted).backward()
        gd = kal._C.render.mesh.dibr_soft_mask_backward_cuda(
            gm.grad, mk, face_idx, prob, cidx, typ, img_s, sigmainv, multiplier)
        assert grad_ok(gd, g32, g64), (rel_err(gd, g64), rel_err(gd, g32))


# ----------------------------------------------------------------- composition + shim
@pytest.mark.parametrize("batch_size", [3, 1])
@pytest.mark.parametrize("flip", [False, True])
def test_dibr_rasterization_composition(kal, batch_size, flip):
    """test_dibr.py:397-529: dibr_rasterization == rasterize(valid=nz>=0) + dibr_soft_mask."""
    g = geo(batch_size, flip)
    z, img, uvs, nz = (g[k] for k in ("face_vertices_z", "face_vertices_image", "face_uvs",
                                      "face_normals_z"))
    feats = [uvs, torch.ones_like(uvs[..., :1])]
    (f_uv, f_one), mask, idx = kal.render.mesh.dibr_rasterization(H, W, z, img, feats, nz)
    (g_uv, g_one), idx2 = kal.render.mesh.rasterize(H, W, z, img, feats, nz >= 0.)
    mask2 = kal.render.mesh.dibr_soft_mask(img, idx2)
    assert torch.equal(idx, idx2) and torch.equal(f_uv, g_uv) and torch.equal(f_one, g_one)
    assert torch.equal(mask, mask2)


def test_raw_ops_match_reference(kal, ref):
    """The _C-shaped shim against the reference's pybind ops, argument for argument."""
    g = geo(3, False)
    z, img, uvs = g["face_vertices_z"], g["face_vertices_image"], g["face_uvs"]
    B, F = z.shape[:2]
    first = torch.arange(B + 1, device=DEV) * F
    img_s = (img * 1000).reshape(-1, 3, 2)
    bbox = torch.cat((img_s.min(dim=1)[0], img_s.max(dim=1)[0]), dim=1)
    args = (H, W, z.reshape(-1, 3), img_s, bbox, uvs.reshape(B * F, 3, -1), first, 1000, 1e-8)
    a = kal._C.render.mesh.packed_rasterize_forward_cuda(*args)
    b = ref.packed_rasterize_forward_cuda(*args)
    for x, y in zip(a, b):
        assert x.dtype == y.dtype and x.shape == y.shape and torch.equal(x, y)
    interp, idx, weights = a
    grad = torch.rand_like(interp)
    ga = kal._C.render.mesh.rasterize_backward_cuda(grad, interp, idx, weights, img, uvs, 1e-8)
    gb = ref.rasterize_backward_cuda(grad.double(), interp.double(), idx, weights.double(),
                                     img.double(), uvs.double(), 1e-8)
    g32 = ref.rasterize_backward_cuda(grad, interp, idx, weights, img, uvs, 1e-8)
    assert grad_ok(ga[0], g32[0], gb[0]) and grad_ok(ga[1], g32[1], gb[1], "grad_face_features")
