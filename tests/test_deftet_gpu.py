"""GPU parity tests of deftet_sparse_render, through the public API and
the _C shim -> ctypes -> C ABI, against

* the reference's own deftet CUDA kernels compiled unmodified into oracle/_ref, wrapped in
  the reference's torch-side code (deftet.py:273-340 restated in ref_forward below),
* the known answers of the reference's tests (tests/deftet_kats.py),
* outputs of the reference's pure-torch naive renderer (tests/golden/deftet_naive.npz).

Bars: face_idx bit-exact; the raw op's depths / weights bit-exact (same arithmetic);
features <= 1e-5; gradients <= 1e-4 relative to the reference run in fp64 on the same
discrete choices (its fp32 atomics are order dependent), or as close to fp64 as the
reference's own fp32 result.  The parametrisation mirrors test_deftet.py:340-437.
"""
import os

import numpy as np
import pytest
import torch

import workloads as wl
import deftet_kats as kats
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


def rel_err(a, b):
    a, b = a.double(), b.double()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def boxes(img):
    return torch.cat((torch.min(img, dim=2)[0], torch.max(img, dim=2)[0]), dim=2)


def ref_forward(ref, pix, rng, z, img, feat, knum, eps=1e-8):
    """DeftetSparseRenderer.forward (deftet.py:273-323) over the reference's own kernel.
    argsort is made stable: the reference leaves the order of equal depths open."""
    B, F = z.shape[:2]
    D = feat.shape[-1]
    P = pix.shape[1]
    face_idx, depth, w0, w1 = ref.deftet_sparse_render_forward_cuda(
        z, img, boxes(img), pix.contiguous(), rng.contiguous(), knum, eps)
    order = torch.argsort(depth, descending=True, dim=-1, stable=True)
    sidx = torch.gather(face_idx, -1, order).contiguous()
    sw0 = torch.gather(w0, -1, order)
    sw1 = torch.gather(w1, -1, order)
    sw2 = (sidx != -1).to(z.dtype) - (sw0 + sw1)
    _idx = (sidx + 1).reshape(B, -1, 1, 1).expand(B, P * knum, 3, D)
    sel = torch.gather(torch.nn.functional.pad(feat, (0, 0, 0, 0, 1, 0), value=0.), 1,
                       _idx).reshape(B, P, knum, 3, D)
    weights = torch.stack([sw0, sw1, sw2], dim=-1).contiguous()
    out = torch.sum(weights.unsqueeze(-1) * sel, dim=-2).contiguous()
    return out, sidx, weights


def check_against_ref(kal, ref, pix, rng, z, img, feat, knum, grads=True):
    dtype = z.dtype
    img1 = img.detach().clone().requires_grad_(True)
    feat1 = feat.detach().clone().requires_grad_(True)
    out, idx = kal.render.mesh.deftet_sparse_render(pix, rng, z, img1, feat1, knum)
    r_out, r_idx, r_w = ref_forward(ref, pix, rng, z, img, feat, knum)
    assert idx.dtype == torch.long and idx.shape == r_idx.shape
    assert torch.equal(idx, r_idx)
    assert out.shape == r_out.shape and out.dtype == dtype
    tol = 1e-5 if dtype == torch.float32 else 1e-12
    assert torch.allclose(out, r_out, rtol=tol, atol=tol)
    if not grads:
        return idx
    g = torch.rand(out.shape, dtype=dtype, device=DEV,
                   generator=torch.Generator(DEV).manual_seed(3))
    out.backward(g)
    r32 = ref.deftet_sparse_render_backward_cuda(g, r_idx, r_w, img, feat, 1e-8)
    r64 = ref.deftet_sparse_render_backward_cuda(g.double(), r_idx, r_w.double(), img.double(),
                                                 feat.double(), 1e-8)
    for mine, a32, a64 in ((img1.grad, r32[0], r64[0]), (feat1.grad, r32[1], r64[1])):
        if float(a64.abs().max()) == 0.:
            assert float(mine.abs().max()) == 0.
            continue
        e64 = rel_err(mine, a64)
        bar = 1e-4 if dtype == torch.float32 else 1e-10
        assert e64 <= bar or e64 <= 1.5 * rel_err(a32, a64) + bar, (e64, rel_err(a32, a64))
    return idx


# ------------------------------------------------------------------ known answers
@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
@pytest.mark.parametrize("cat_features", [False, True])
@pytest.mark.parametrize("case,lo", [("full", -4.), ("restricted", -2.1)])
def test_reference_kats(kal, dtype, cat_features, case, lo):
    pix, z, img, feats = (kats.inputs(dtype)[i] for i in range(4))
    pix, z, img = pix.to(DEV), z.to(DEV), img.to(DEV)
    feats = [f.to(DEV) for f in feats]
    rng = kats.ranges(lo, dtype).to(DEV)
    gt_idx, gt_f0, gt_f1 = (t.to(DEV) for t in kats.expected(case, dtype))
    if cat_features:
        out, idx = kal.render.mesh.deftet_sparse_render(pix, rng, z, img,
                                                        torch.cat(feats, dim=-1), 5)
        assert torch.equal(idx, gt_idx)
        assert torch.allclose(out, torch.cat([gt_f0, gt_f1], dim=-1), atol=3e-3, rtol=1e-5)
    else:
        out, idx = kal.render.mesh.deftet_sparse_render(pix, rng, z, img, feats, 5)
        assert isinstance(out, tuple) and len(out) == 2
        assert torch.equal(idx, gt_idx)
        assert torch.allclose(out[0], gt_f0)
        assert torch.allclose(out[1], gt_f1, atol=3e-3, rtol=1e-5)


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
@pytest.mark.parametrize("centre", [0, 1])
def test_naive_golden(kal, dtype, centre):
    g = np.load(os.path.join(wl.GOLDEN, "deftet_naive.npz"))
    tag = str(dtype)[-7:]
    key = "%s_centre%d_k20" % (tag, centre)
    geo = wl.fixture_geometry(batch_size=3, dtype=dtype)
    z, img, uvs = (geo[k].to(DEV) for k in ("face_vertices_z", "face_vertices_image",
                                            "face_uvs"))
    img.requires_grad_(True)
    uvs.requires_grad_(True)
    pix = torch.from_numpy(g[tag + "_pixel_coords"]).to(DEV)
    rr = torch.from_numpy(g[key + "_ranges"]).to(DEV)
    out, idx = kal.render.mesh.deftet_sparse_render(pix, rr, z, img, uvs, 20)
    assert torch.equal(idx.cpu(), torch.from_numpy(g[key + "_idx"]).long())
    assert torch.allclose(out.cpu(), torch.from_numpy(g[key + "_feat"]), rtol=1e-4, atol=1e-4)
    out.backward(torch.from_numpy(g[key + "_grad_out"]).to(DEV))
    assert torch.allclose(img.grad.cpu(), torch.from_numpy(g[key + "_grad_img"]),
                          rtol=5e-3, atol=5e-3)
    assert torch.allclose(uvs.grad.cpu(), torch.from_numpy(g[key + "_grad_feat"]),
                          rtol=1e-3, atol=1e-3)


# ------------------------------------------------------------------ reference test matrix
@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
@pytest.mark.parametrize("batch_size", [1, 3])
@pytest.mark.parametrize("num_pixels", [1, 31, 1025])
@pytest.mark.parametrize("render_up_to_center", [True, False])
@pytest.mark.parametrize("knum", [20, 30])
def test_model_obj_vs_ref(kal, ref, dtype, batch_size, num_pixels, render_up_to_center, knum):
    geo = wl.fixture_geometry(batch_size=batch_size, dtype=dtype)
    z, img, uvs = (geo[k].to(DEV) for k in ("face_vertices_z", "face_vertices_image",
                                            "face_uvs"))
    vc = geo["vertices_camera"].to(DEV)
    gen = torch.Generator(DEV).manual_seed(num_pixels + knum)
    pix = torch.rand((batch_size, num_pixels, 2), device=DEV, dtype=dtype, generator=gen) * 2 - 1
    min_z, max_z = vc[:, :, -1].min(dim=1)[0], vc[:, :, -1].max(dim=1)[0]
    lo = (min_z + max_z) / 2. if render_up_to_center else min_z
    rng = torch.nn.functional.pad(lo.unsqueeze(-1), (0, 1), value=0.).unsqueeze(1).repeat(
        1, num_pixels, 1)
    feat = torch.cat([uvs, torch.ones_like(uvs[..., :1])], dim=-1)
    check_against_ref(kal, ref, pix, rng, z, img, feat, knum)


def test_autograd_contract(kal):
    """No gradient for pixel_coords, render_ranges, face_vertices_z (deftet.py:340,
    test_deftet.py:479-481)."""
    geo = wl.fixture_geometry(batch_size=1)
    z, img, uvs = (geo[k].to(DEV).requires_grad_(True)
                   for k in ("face_vertices_z", "face_vertices_image", "face_uvs"))
    pix = (torch.rand((1, 64, 2), device=DEV) * 2 - 1).requires_grad_(True)
    rng = torch.tensor([[[-10., 0.]]], device=DEV).repeat(1, 64, 1).requires_grad_(True)
    out, idx = kal.render.mesh.deftet_sparse_render(pix, rng, z, img, uvs, 8)
    assert not idx.requires_grad
    out.sum().backward()
    assert pix.grad is None and rng.grad is None and z.grad is None
    assert img.grad is not None and uvs.grad is not None


# ------------------------------------------------------------------ deep stacks, overflow
def random_soup(F, B, dtype, seed, size=0.6):
    """F random triangles per view scattered over [-1, 1]^2: every point is under many."""
    g = torch.Generator().manual_seed(seed)
    c = torch.rand((B, F, 1, 2), generator=g, dtype=dtype) * 2 - 1
    img = c + (torch.rand((B, F, 3, 2), generator=g, dtype=dtype) - 0.5) * size
    z = -1 - 3 * torch.rand((B, F, 3), generator=g, dtype=dtype)
    feat = torch.rand((B, F, 3, 4), generator=g, dtype=dtype)
    return z.to(DEV), img.to(DEV), feat.to(DEV)


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
@pytest.mark.parametrize("F,size,knum", [(12000, 0.6, 300), (3000, 0.6, 10), (3000, 0.6, 1),
                                         (20000, 0.05, 30), (500, 3.0, 8), (5000, 0.2, 33),
                                         (12000, 0.6, 40), (40000, 0.6, 300)])
def test_soup_vs_ref(kal, ref, dtype, F, size, knum):
    """Hundreds of layers per point (rank sort up to 128 hits, bitonic network beyond), more hits than knum
    (the first knum by face index must survive), faces larger than the sample domain (wide
    list), tiny faces (fine grid)."""
    B, P = 2, 777
    z, img, feat = random_soup(F, B, dtype, seed=F + knum, size=size)
    gen = torch.Generator(DEV).manual_seed(1)
    pix = torch.rand((B, P, 2), device=DEV, dtype=dtype, generator=gen) * 2 - 1
    rng = torch.tensor([[[-3.5, -1.2]]], device=DEV, dtype=dtype).repeat(B, P, 1)
    idx = check_against_ref(kal, ref, pix, rng, z, img, feat, knum)
    hits = (idx >= 0).sum(-1)
    if knum in (10, 1, 8, 40):
        assert int(hits.max()) == knum        # the cut is exercised
    if knum == 300:
        assert int(hits.max()) > (128 if F == 40000 else 32)   # bitonic network / rank sort


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_raw_ops_vs_ref(kal, ref, dtype):
    """The _C shim's raw ops (bindings.cpp:82-83): same return lists, bit-exact slots."""
    B, P, F, knum = 2, 500, 4000, 12
    z, img, feat = random_soup(F, B, dtype, seed=9, size=0.4)
    gen = torch.Generator(DEV).manual_seed(2)
    pix = torch.rand((B, P, 2), device=DEV, dtype=dtype, generator=gen) * 2.4 - 1.2
    rng = torch.tensor([[[-3.9, -1.1]]], device=DEV, dtype=dtype).repeat(B, P, 1)
    bb = boxes(img)
    mine = kal._C.render.mesh.deftet_sparse_render_forward_cuda(z, img, bb, pix, rng, knum, 1e-8)
    theirs = ref.deftet_sparse_render_forward_cuda(z, img, bb, pix, rng, knum, 1e-8)
    assert len(mine) == 4
    for a, b in zip(mine, theirs):
        assert a.dtype == b.dtype and torch.equal(a, b)
    assert int((mine[0] >= 0).sum(-1).max()) == knum
    # backward raw op on the reference's own sorted state
    out, sidx, w = ref_forward(ref, pix, rng, z, img, feat, knum)
    g = torch.rand_like(out)
    gi, gf = kal._C.render.mesh.deftet_sparse_render_backward_cuda(g, sidx, w, img, feat, 1e-8)
    r64 = ref.deftet_sparse_render_backward_cuda(g.double(), sidx, w.double(), img.double(),
                                                 feat.double(), 1e-8)
    r32 = ref.deftet_sparse_render_backward_cuda(g, sidx, w, img, feat, 1e-8)
    bar = 1e-4 if dtype == torch.float32 else 1e-10
    for mine_g, a32, a64 in ((gi, r32[0], r64[0]), (gf, r32[1], r64[1])):
        e = rel_err(mine_g, a64)
        assert e <= bar or e <= 1.5 * rel_err(a32, a64) + bar, e


def test_caller_boxes_are_honoured(kal, ref):
    """The raw op takes face_bboxes from the caller (deftet.cpp:47-53); shrunken boxes must
    cut hits exactly as in the reference."""
    B, P, F, knum = 1, 400, 1500, 16
    z, img, _ = random_soup(F, B, torch.float32, seed=4, size=0.5)
    pix = torch.rand((B, P, 2), device=DEV, generator=torch.Generator(DEV).manual_seed(8)) * 2 - 1
    rng = torch.tensor([[[-9., 0.]]], device=DEV).repeat(B, P, 1)
    bb = boxes(img)
    mid = (bb[..., :2] + bb[..., 2:]) / 2
    bb = torch.cat((mid - (mid - bb[..., :2]) * 0.5, mid + (bb[..., 2:] - mid) * 0.5), dim=-1)
    mine = kal._C.render.mesh.deftet_sparse_render_forward_cuda(z, img, bb, pix, rng, knum, 1e-8)
    theirs = ref.deftet_sparse_render_forward_cuda(z, img, bb, pix, rng, knum, 1e-8)
    for a, b in zip(mine, theirs):
        assert torch.equal(a, b)


# ------------------------------------------------------------------ edges
def test_edge_cases(kal, ref):
    dt = torch.float32
    z, img, feat = random_soup(200, 2, dt, seed=1, size=0.8)
    # no sample points / knum 0 / no faces
    pix0 = torch.empty((2, 0, 2), device=DEV)
    out, idx = kal.render.mesh.deftet_sparse_render(pix0, pix0.clone(), z, img, feat, 5)
    assert out.shape == (2, 0, 5, 4) and idx.shape == (2, 0, 5)
    pix = torch.rand((2, 33, 2), device=DEV) * 2 - 1
    rng = torch.tensor([[[-9., 0.]]], device=DEV).repeat(2, 33, 1)
    out, idx = kal.render.mesh.deftet_sparse_render(pix, rng, z, img, feat, 0)
    assert out.shape == (2, 33, 0, 4) and idx.shape == (2, 33, 0)
    out, idx = kal.render.mesh.deftet_sparse_render(pix, rng, z[:, :0], img[:, :0], feat[:, :0], 4)
    assert bool((idx == -1).all()) and bool((out == 0).all())
    # NaN / inf coordinates in samples and faces, a degenerate (single point) sample set,
    # identical samples, zero-area faces
    pix2 = pix.clone()
    pix2[0, 3] = float("nan")
    pix2[1, 5, 0] = float("inf")
    pix2[1, 7] = pix2[1, 8]
    img2 = img.clone()
    img2[0, 10, 1, 0] = float("nan")
    img2[1, 11, 2, 1] = float("inf")
    img2[1, 12, 1] = img2[1, 12, 0]
    check_against_ref(kal, ref, pix2, rng, z, img2, feat, 12, grads=False)
    one = pix[:, :1].repeat(1, 9, 1)
    check_against_ref(kal, ref, one, rng[:, :9], z, img, feat, 12)
    # samples far outside the mesh, empty depth range
    check_against_ref(kal, ref, pix + 5., rng, z, img, feat, 3)
    check_against_ref(kal, ref, pix, rng * 0., z, img, feat, 3)


def test_errors(kal):
    z, img, feat = random_soup(10, 1, torch.float32, seed=1)
    pix = torch.zeros((1, 4, 2), device=DEV)
    with pytest.raises(RuntimeError):
        kal.render.mesh.deftet_sparse_render(pix.cpu(), pix.cpu(), z.cpu(), img.cpu(),
                                             feat.cpu(), 3)
    with pytest.raises(TypeError):
        kal.render.mesh.deftet_sparse_render(pix.double(), pix, z, img, feat, 3)
    with pytest.raises(RuntimeError):
        kal.render.mesh.deftet_sparse_render(pix, pix, z, img[:, :5], feat, 3)
    with pytest.raises(ValueError):     # staging of knum hits exceeds an SM's shared memory
        kal.render.mesh.deftet_sparse_render(pix, pix, z, img, feat, 20000)


def test_deterministic(kal):
    z, img, feat = random_soup(3000, 2, torch.float32, seed=2, size=0.5)
    pix = torch.rand((2, 999, 2), device=DEV) * 2 - 1
    rng = torch.tensor([[[-9., 0.]]], device=DEV).repeat(2, 999, 1)
    a = kal.render.mesh.deftet_sparse_render(pix, rng, z, img, feat, 17)
    b = kal.render.mesh.deftet_sparse_render(pix, rng, z, img, feat, 17)
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
