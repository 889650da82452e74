"""prepare_vertices / index_vertices_by_faces / scatter_mesh_grads on the B200 against the
mesh oracle and the golden vectors of the reference's own functions.
Gathers are bit-exact; normals and gradients within fp32 tolerance (1e-5 relative)."""
import pytest
import torch

import kaolin_windows_b200 as kal
from oracle import mesh_oracle as mo
from test_mesh_golden import compare, load, run

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.mark.parametrize("mode", ["rot_trans", "transform"])
def test_prepare_vertices_matches_reference_golden(mode):
    d = load()
    res, grads = run(kal.render.mesh.prepare_vertices, d, mode, DEV)
    compare(res, grads, d, mode, 1e-5)


@pytest.mark.parametrize("B,V,F", [(1, 5, 1), (3, 100, 517), (2, 2000, 40000)])
def test_face_setup_gathers_are_bit_exact_and_grads_match(B, V, F):
    from kaolin_windows_b200.render.mesh.utils import _FaceSetup
    g = torch.Generator(device=DEV).manual_seed(B * 1000 + F)
    vc = torch.randn((B, V, 3), device=DEV, generator=g)
    vi = torch.randn((B, V, 2), device=DEV, generator=g)
    faces = torch.randint(0, V, (F, 3), device=DEV, generator=g)
    if F > 1:   # no repeated corners (zero-area faces have 1e10-sized normal gradients)
        ok = (faces[:, 0] != faces[:, 1]) & (faces[:, 1] != faces[:, 2]) & (faces[:, 0] != faces[:, 2])
        faces = faces[ok].contiguous()
        F = faces.shape[0]
    cot = [torch.randn(s, device=DEV, generator=g) for s in ((B, F, 3, 3), (B, F, 3, 2), (B, F, 3))]
    a, b = vc.clone().requires_grad_(True), vi.clone().requires_grad_(True)
    fvc, fvi, fn = _FaceSetup.apply(a, b, faces)
    torch.autograd.backward((fvc, fvi, fn), cot)

    def chain(dtype):
        a2, b2 = vc.to(dtype).requires_grad_(True), vi.to(dtype).requires_grad_(True)
        wvc = mo.index_vertices_by_faces(a2, faces)
        wvi = mo.index_vertices_by_faces(b2, faces)
        wn = mo.face_normals(wvc, unit=True)
        torch.autograd.backward((wvc, wvi, wn), [c.to(dtype) for c in cot])
        return wvc, wvi, wn, a2.grad, b2.grad

    wvc, wvi, wn, ga, gb = chain(torch.float64)
    _, _, n32, ga32, gb32 = chain(torch.float32)      # the reference's own fp32 (torch kernels)
    assert torch.equal(fvc, wvc.float()) and torch.equal(fvi, wvi.float())
    # fp32 bar: 1e-5 of the largest value, or as close to fp64 as torch's fp32 (thin faces)
    for got, want, ref32 in ((fn, wn, n32), (a.grad, ga, ga32), (b.grad, gb, gb32)):
        err = float((got.detach().double() - want.detach()).abs().max())
        err32 = float((ref32.detach().double() - want.detach()).abs().max())
        assert err <= max(1e-5 * float(want.abs().max()), 2 * err32), (err, err32)


def test_face_setup_partial_cotangents():
    d = load()
    v = d["vertices"].to(DEV).requires_grad_(True)
    kw = dict(camera_rot=d["camera_rot"].to(DEV), camera_trans=d["camera_trans"].to(DEV))
    faces = d["faces"].long().to(DEV)
    proj = d["camera_proj"].to(DEV)
    for use in ((0,), (1,), (2,), (0, 2)):
        res = kal.render.mesh.prepare_vertices(v, faces, proj, **kw)
        got, = torch.autograd.grad([res[i] for i in use], v, [d["cot%d" % i].to(DEV) for i in use])
        def oracle(dtype, dev):
            v2 = d["vertices"].to(dev, dtype).requires_grad_(True)
            ref = mo.prepare_vertices(v2, faces.to(dev), proj.to(dev, dtype),
                                      **{k: t.to(dev, dtype) for k, t in kw.items()})
            w, = torch.autograd.grad([ref[i] for i in use], v2,
                                     [d["cot%d" % i].to(dev, dtype) for i in use])
            return w.cpu().double()

        want, w32 = oracle(torch.float64, "cpu"), oracle(torch.float32, DEV)
        err = float((got.detach().cpu().double() - want).abs().max())
        err32 = float((w32 - want).abs().max())     # torch's own fp32 kernels, same chain
        assert err <= max(1e-5 * float(want.abs().max()), 2 * err32), (use, err, err32)


@pytest.mark.parametrize("B,V,F,C", [(1, 4, 1, 1), (2, 300, 1000, 3), (3, 50, 777, 7), (1, 10, 0, 3)])
def test_index_vertices_by_faces(B, V, F, C):
    g = torch.Generator(device=DEV).manual_seed(V + F + C)
    vf = torch.randn((B, V, C), device=DEV, generator=g)
    faces = torch.randint(0, V, (F, 3), device=DEV, generator=g)
    x = vf.clone().requires_grad_(True)
    out = kal.ops.mesh.index_vertices_by_faces(x, faces)
    want = mo.index_vertices_by_faces(vf, faces)
    assert out.shape == (B, F, 3, C) and torch.equal(out, want)
    cot = torch.randn(out.shape, device=DEV, generator=g)
    out.backward(cot)
    wg = torch.zeros((B, V, C), device=DEV, dtype=torch.float64)
    wg.index_add_(1, faces.reshape(-1), cot.double().reshape(B, F * 3, C))
    assert torch.allclose(x.grad.double(), wg, rtol=1e-5, atol=1e-5)


def test_index_vertices_by_faces_errors():
    vf = torch.zeros((1, 4, 3), device=DEV)
    faces = torch.zeros((2, 3), dtype=torch.long, device=DEV)
    with pytest.raises(AssertionError):
        kal.ops.mesh.index_vertices_by_faces(vf[0], faces)
    with pytest.raises(AssertionError):
        kal.ops.mesh.index_vertices_by_faces(vf, faces[0])
    with pytest.raises(RuntimeError):
        kal.ops.mesh.index_vertices_by_faces(vf.cpu(), faces.cpu())
    with pytest.raises(TypeError):
        kal.ops.mesh.index_vertices_by_faces(vf.double(), faces)


def test_scatter_mesh_grads_is_view_summed_index_add():
    B, V, F, D = 4, 500, 3000, 3
    g = torch.Generator(device=DEV).manual_seed(3)
    faces = torch.randint(0, V, (F, 3), device=DEV, generator=g)
    gi = torch.randn((B, F, 3, 2), device=DEV, generator=g)
    gf = torch.randn((B, F, 3, D), device=DEV, generator=g)
    flat = kal.parallel.scatter_mesh_grads(faces, V, gi, gf)
    want = torch.zeros((V, 2 + D), device=DEV, dtype=torch.float64)
    want.index_add_(0, faces.reshape(-1),
                    torch.cat([gi, gf], dim=-1).double().sum(0).reshape(-1, 2 + D))
    assert flat.shape == (V, 2 + D)
    assert torch.allclose(flat.double(), want, rtol=1e-5, atol=1e-5)


def test_prepare_vertices_feeds_the_path():
    """vertices -> prepare_vertices -> dibr_rasterization -> backward down to the vertices,
    against the same chain through the oracle's prepare_vertices (torch) and the B200 path."""
    d = load()
    faces = d["faces"].long().to(DEV)
    proj = d["camera_proj"].to(DEV)
    kw = dict(camera_rot=d["camera_rot"].to(DEV), camera_trans=d["camera_trans"].to(DEV))
    B, V = d["vertices"].shape[:2]
    colors = torch.rand((B, V, 3), device=DEV, generator=torch.Generator(device=DEV).manual_seed(1))
    grads = []
    for prep, gather in ((kal.render.mesh.prepare_vertices, kal.ops.mesh.index_vertices_by_faces),
                         (mo.prepare_vertices, mo.index_vertices_by_faces)):
        v = d["vertices"].to(DEV).requires_grad_(True)
        c = colors.clone().requires_grad_(True)
        fvc, fvi, fn = prep(v, faces, proj, **kw)
        out, mask, idx = kal.render.mesh.dibr_rasterization(
            64, 64, fvc[..., 2].contiguous(), fvi.contiguous(), gather(c, faces).contiguous(),
            fn[..., 2].contiguous())
        (out.sum() + mask.sum()).backward()
        grads.append((v.grad, c.grad, idx))
    assert torch.equal(grads[0][2], grads[1][2])
    for a, b in zip(grads[0][:2], grads[1][:2]):
        assert float((a - b).abs().max()) <= 1e-4 * float(b.abs().max())


def test_mesh_grad_exchange_graph_matches_eager():
    B, V, F, D = 3, 400, 2500, 3
    g = torch.Generator(device=DEV).manual_seed(9)
    faces = torch.randint(0, V, (F, 3), device=DEV, generator=g)
    gi = torch.randn((B, F, 3, 2), device=DEV, generator=g)
    gf = torch.randn((B, F, 3, D), device=DEV, generator=g)
    ex = kal.parallel.MeshGradExchange(faces, V)
    for it in range(4):                      # eager, capture, replay, replay on new data
        if it == 3:
            gi.mul_(-0.5)
        want = kal.parallel.scatter_mesh_grads(faces, V, gi, gf)
        got = ex(gi, gf)
        assert torch.allclose(got, want, rtol=1e-5, atol=1e-5)
    assert len(ex.graphs) == 1
