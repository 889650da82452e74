"""Pins the CPU oracle (oracle/dibr_oracle.cpp) against the reference's own golden
vectors, re-packed under tests/golden/ by tests/golden/make_golden.py.

Mirrors tests/python/kaolin/render/mesh/test_dibr.py (TestSimpleDibrSoftMask,
TestDibrSoftMask, TestDibrRasterization) and test_rasterization.py (TestRasterize) of
the reference, with the same parametrisation and the same tolerances, but running the
oracle on CPU instead of the CUDA op.
"""
import os

import numpy as np
import pytest
import torch

import workloads as wl
from oracle import dibr_oracle as orc

GOLDEN = wl.GOLDEN
H, W = 35, 31


def _g(npz, name, dtype=None):
    t = torch.from_numpy(np.asarray(npz[name]))
    return t.to(dtype) if dtype is not None else t


@pytest.fixture(scope="module")
def simple_gt():
    return np.load(os.path.join(GOLDEN, "dibr_simple.npz"))


@pytest.fixture(scope="module")
def sphere_gt():
    return np.load(os.path.join(GOLDEN, "dibr_sphere.npz"))


@pytest.fixture(scope="module")
def naive_gt():
    return np.load(os.path.join(GOLDEN, "rasterize_naive.npz"))


def _large_bboxes(img_scaled, boxlen, multiplier):
    pmin = torch.min(img_scaled, dim=-2)[0]
    pmax = torch.max(img_scaled, dim=-2)[0]
    return torch.cat([pmin - boxlen * multiplier, pmax + boxlen * multiplier], dim=-1)


# ------------------------------------------------------------------ simple triangles
@pytest.mark.parametrize("dtype", [torch.float, torch.double])
@pytest.mark.parametrize("sigmainv", [7000, 70])
@pytest.mark.parametrize("boxlen", [0.02, 0.2])
class TestSimpleDibrSoftMask:
    @pytest.fixture
    def inputs(self, dtype):
        img, z = wl.simple_soft_mask_fixture(dtype)
        feat = torch.zeros(z.shape + (1,), dtype=dtype)
        _, face_idx, _ = orc.rasterize_forward(H, W, z, img, feat)
        return img, face_idx

    @pytest.mark.parametrize("multiplier", [1000, 100, 1])
    @pytest.mark.parametrize("knum", [30, 20])
    def test_C_forward(self, inputs, simple_gt, dtype, sigmainv, boxlen, knum, multiplier):
        img, face_idx = inputs
        key = "%d_%d_%d_%s" % (H, W, sigmainv, boxlen)
        img_s = img * multiplier
        mask, prob, cidx, typ = orc.dibr_soft_mask_forward(
            img_s, _large_bboxes(img_s, boxlen, multiplier), face_idx, sigmainv, knum,
            multiplier)
        assert torch.allclose(mask, _g(simple_gt, "soft_mask_" + key, dtype),
                              atol=1e-5, rtol=1e-5)
        assert torch.equal(cidx, _g(simple_gt, "close_face_idx_" + key).long()[..., :knum])
        assert torch.allclose(prob, _g(simple_gt, "close_face_dist_" + key, dtype)[..., :knum],
                              atol=1e-5, rtol=1e-5)
        assert torch.equal(typ, _g(simple_gt, "close_face_dist_type_" + key)[..., :knum])

    @pytest.mark.parametrize("multiplier", [1000, 100, 1])
    @pytest.mark.parametrize("knum", [30, 20])
    def test_backward(self, inputs, simple_gt, dtype, sigmainv, boxlen, knum, multiplier):
        img, face_idx = inputs
        key = "%d_%d_%d_%s" % (H, W, sigmainv, boxlen)
        mask, ctx = orc.dibr_soft_mask_forward_py(img, face_idx, sigmainv, boxlen, knum,
                                                  multiplier)
        m = mask.detach().clone().requires_grad_(True)
        shifted = torch.nn.functional.pad(face_idx != -1, (0, 5))[..., 5:]
        orc.mask_iou(m, shifted).backward()
        grad = orc.dibr_soft_mask_backward_from_ctx(ctx, m.grad)
        assert torch.allclose(grad, _g(simple_gt, "grad_face_vertices_image_" + key, dtype),
                              rtol=1e-5, atol=1e-5)


# ------------------------------------------------------------------------- model.obj
@pytest.mark.parametrize("dtype", [torch.float, torch.double])
@pytest.mark.parametrize("batch_size", [1, 3])
@pytest.mark.parametrize("flip", [False, True])
@pytest.mark.parametrize("sigmainv", [7000, 70])
@pytest.mark.parametrize("boxlen", [0.02, 0.01])
class TestDibrSoftMask:
    @pytest.fixture
    def inputs(self, dtype, batch_size, flip):
        geo = wl.fixture_geometry(batch_size=batch_size, flip=flip, dtype=dtype)
        z, img = geo["face_vertices_z"], geo["face_vertices_image"]
        feat = torch.zeros(z.shape + (1,), dtype=dtype)
        _, face_idx, _ = orc.rasterize_forward(H, W, z, img, feat)
        return img, face_idx

    @pytest.mark.parametrize("multiplier", [1000, 100])
    @pytest.mark.parametrize("knum", [30, 40])
    def test_C_forward(self, inputs, sphere_gt, dtype, batch_size, sigmainv, boxlen, knum,
                       multiplier):
        img, face_idx = inputs
        key = "%d_%d_%d_%s" % (H, W, sigmainv, boxlen)
        img_s = img * multiplier
        mask, prob, cidx, typ = orc.dibr_soft_mask_forward(
            img_s, _large_bboxes(img_s, boxlen, multiplier), face_idx, sigmainv, knum,
            multiplier)
        b = batch_size
        assert torch.allclose(mask, _g(sphere_gt, "soft_mask_" + key, dtype)[:b],
                              atol=1e-5, rtol=1e-5)
        assert torch.equal(cidx, _g(sphere_gt, "close_face_idx_" + key).long()[:b, ..., :knum])
        assert torch.allclose(prob, _g(sphere_gt, "close_face_dist_" + key, dtype)[:b, ..., :knum],
                              atol=1e-5, rtol=1e-5)
        diff = typ != _g(sphere_gt, "close_face_dist_type_" + key)[:b, ..., :knum]
        assert torch.sum(diff) / diff.numel() <= 0.01       # test_dibr.py:340-341

    @pytest.mark.parametrize("multiplier", [1000, 100, 1])
    @pytest.mark.parametrize("knum", [30, 40])
    def test_backward(self, inputs, sphere_gt, dtype, batch_size, sigmainv, boxlen, knum,
                      multiplier):
        img, face_idx = inputs
        key = "%d_%d_%d_%s" % (H, W, sigmainv, boxlen)
        mask, ctx = orc.dibr_soft_mask_forward_py(img, face_idx, sigmainv, boxlen, knum,
                                                  multiplier)
        m = mask.detach().clone().requires_grad_(True)
        shifted = torch.nn.functional.pad(face_idx != -1, (0, 5))[..., 5:]
        orc.mask_iou(m, shifted).backward()
        grad = orc.dibr_soft_mask_backward_from_ctx(ctx, m.grad)
        gt = _g(sphere_gt, "grad_face_vertices_image_" + key, dtype)[:batch_size]
        assert torch.allclose(grad, gt, rtol=1e-1, atol=1e-1)  # test_dibr.py:390-394


# --------------------------------------------------- rasterize vs the naive torch oracle
@pytest.mark.parametrize("dtype", [torch.float, torch.double])
@pytest.mark.parametrize("batch_size", [1, 3])
@pytest.mark.parametrize("flip", [False, True])
@pytest.mark.parametrize("with_valid_faces", [False, True])
class TestRasterize:
    def _run(self, naive_gt, dtype, batch_size, flip, with_valid_faces):
        geo = wl.fixture_geometry(batch_size=batch_size, flip=flip, dtype=dtype)
        z, img, uvs = geo["face_vertices_z"], geo["face_vertices_image"], geo["face_uvs"]
        valid = None
        dn = str(dtype)[-7:]
        if with_valid_faces:  # test_rasterization.py:91-96, stored with the golden
            valid = _g(naive_gt, "%s_flip%d_validmask" % (dn, flip))[:batch_size]
        return z, img, uvs, valid, "%s_flip%d_valid%d" % (dn, flip, with_valid_faces)

    def test_forward_with_list(self, naive_gt, dtype, batch_size, flip, with_valid_faces):
        z, img, uvs, valid, key = self._run(naive_gt, dtype, batch_size, flip, with_valid_faces)
        feats = [uvs, torch.ones_like(uvs[..., 1:])]
        interp, face_idx, _ = orc.rasterize_forward(H, W, z, img, feats, valid)
        assert torch.equal(face_idx, _g(naive_gt, key + "_idx").long()[:batch_size])
        assert torch.allclose(interp, _g(naive_gt, key + "_feat", dtype)[:batch_size],
                              rtol=1e-5, atol=1e-5)

    def test_backward_with_list(self, naive_gt, dtype, batch_size, flip, with_valid_faces):
        z, img, uvs, valid, key = self._run(naive_gt, dtype, batch_size, flip, with_valid_faces)
        feats = [uvs, torch.ones_like(uvs[..., 1:])]
        interp, face_idx, ctx = orc.rasterize_forward(H, W, z, img, feats, valid)
        grad_out = _g(naive_gt, key + "_grad_out", dtype)[:batch_size].contiguous()
        gi, gf = orc.rasterize_backward_from_ctx(ctx, grad_out)
        # tolerances of test_rasterization.py:232-236
        assert torch.allclose(gi, _g(naive_gt, key + "_grad_img", dtype)[:batch_size],
                              rtol=1e-3, atol=1e-2)
        assert torch.allclose(gf, _g(naive_gt, key + "_grad_feat", dtype)[:batch_size],
                              rtol=1e-3, atol=1e-3)


def test_dibr_rasterization_composition():
    """test_dibr.py:397-529: dibr_rasterization == rasterize(valid = nz >= 0) + soft mask."""
    geo = wl.fixture_geometry(batch_size=3)
    z, img, uvs, nz = (geo[k] for k in ("face_vertices_z", "face_vertices_image", "face_uvs",
                                         "face_normals_z"))
    interp, mask, idx, _ = orc.dibr_rasterization_forward(H, W, z, img, uvs, nz)
    i2, idx2, _ = orc.rasterize_forward(H, W, z, img, uvs, nz >= 0.)
    m2, _ = orc.dibr_soft_mask_forward_py(img, idx2)
    assert torch.equal(idx, idx2) and torch.equal(interp, i2) and torch.equal(mask, m2)
    assert (idx >= 0).sum() > 100 and ((mask > 0) & (mask < 1)).sum() > 10


@pytest.mark.parametrize("H,W", [(35, 31), (64, 64), (200, 333)])
def test_closed_form_pair_count_matches_oracle_counter(H, W):
    """bench.py's secondary roofline counts N_pair with
    workloads.raster_pair_count; the oracle counts the same pairs while it rasterizes."""
    g = wl.fixture_geometry(batch_size=3)
    z, img, feat, nz = (g[k] for k in ("face_vertices_z", "face_vertices_image", "face_uvs",
                                       "face_normals_z"))
    _, _, ctx = orc.rasterize_forward(H, W, z, img, feat, nz >= 0, return_pair_count=True)
    assert ctx["pair_count"] == wl.raster_pair_count(H, W, img, nz >= 0) > 0


def test_exact_weights64_is_the_fp64_forward_of_the_oracle():
    """tests/parity_util.exact_weights64 (the weights of the low-noise gradient reference)
    against the weights the oracle's forward kernel saves when instantiated for double."""
    import parity_util as pu
    g = wl.fixture_geometry(batch_size=2)
    z, img, feat, nz = (g[k] for k in ("face_vertices_z", "face_vertices_image", "face_uvs",
                                       "face_normals_z"))
    for mult, eps in ((1000, 1e-8), (1, 1e-8), (100, 1e-6)):
        _, idx, ctx = orc.rasterize_forward(64, 48, z.double(), img.double(), feat.double(),
                                            nz >= 0, multiplier=mult, eps=eps)
        w = pu.exact_weights64(64, 48, img, idx, mult, eps)
        assert (idx >= 0).sum() > 200
        assert float((w - ctx["weights"]).abs().max()) <= 1e-12
