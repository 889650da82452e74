"""dibr_rasterization_mesh (the renderer reading vertices through
`faces`) against the face-tensor path and against the reference.

* forward: bit-identical (torch.equal) to dibr_rasterization on the face tensors
  prepare_vertices / index_vertices_by_faces build from the same arrays -- the chain the
  reference runs (kaolin/render/mesh/utils.py:128-175, ops/mesh/mesh.py:25-47,
  render/mesh/dibr.py:119-209);
* backward: vertex-level gradients against the reference's backward kernels run in fp64
  (tests/parity_util.py) and scattered to the vertices in double -- the bar per VERTEX is
  the per-face bar of parity_util (1e-4 of the vertex's gradient, floored at 1e-3 of the
  largest).
"""
import pytest
import torch

import parity_util as pu
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


def face_tensors(kal, inp):
    """The reference's chain on our own kernels: prepare_vertices' gather / normal step and
    index_vertices_by_faces."""
    vcam, vimg, faces, vfeat = (inp[k] for k in ("vertices_camera", "vertices_image", "faces",
                                                 "vertex_features"))
    fvc, fvi, fn = kal.render.mesh.utils._FaceSetup.apply(vcam, vimg, faces)
    feat = kal.ops.mesh.index_vertices_by_faces(vfeat.contiguous(), faces)
    return fvc[..., 2].contiguous(), fvi, feat, fn[..., 2].contiguous()


def scatter64(face_grad, faces, V):
    """(B, F, 3, C) face-corner gradients -> (B, V, C) in double."""
    B, F, _, C = face_grad.shape
    out = torch.zeros((B, V, C), dtype=torch.float64, device=face_grad.device)
    out.index_add_(1, faces.reshape(-1), face_grad.double().reshape(B, F * 3, C))
    return out


def vertex_check(name, mine, ref32, ref64):
    # parity_util works on (B, N, 3, C): one "face" per vertex
    st = pu.face_grad_check(name, mine.unsqueeze(2), ref32.unsqueeze(2), ref64.unsqueeze(2))
    return st


def compare(kal, ref, inp, H, W, label, shared=False, grads=True, **kw):
    vcam, vimg, faces, vfeat = (inp[k] for k in ("vertices_camera", "vertices_image", "faces",
                                                 "vertex_features"))
    vfeat = vfeat.contiguous()
    if shared:
        vfeat = vfeat[:1].contiguous()
    B, V = vcam.shape[:2]
    z, img, feat, nz = face_tensors(kal, dict(inp, vertex_features=vfeat.expand(B, -1, -1).contiguous()))
    o_f, m_f, i_f = kal.render.mesh.dibr_rasterization(H, W, z, img, feat, nz, **kw)
    vimg_l = vimg.clone().requires_grad_(grads)
    vfeat_l = vfeat.clone().requires_grad_(grads)
    o_m, m_m, i_m = kal.render.mesh.dibr_rasterization_mesh(H, W, vcam, vimg_l, faces, vfeat_l,
                                                            **kw)
    assert torch.equal(i_m, i_f), "face index mismatch: %d px" % int((i_m != i_f).sum())
    assert torch.equal(o_m, o_f) and torch.equal(m_m, m_f)
    stats = dict(label=label, H=H, W=W, shape=[B, int(faces.shape[0]), V], params=kw,
                 covered=int((i_m >= 0).sum()), soft=int(((m_m > 0) & (m_m < 1)).sum()))
    if grads:
        gen = torch.Generator(device=DEV).manual_seed(3)
        g1 = torch.rand(o_m.shape, device=DEV, generator=gen)
        g2 = torch.rand(m_m.shape, device=DEV, generator=gen)
        torch.autograd.backward((o_m, m_m), (g1, g2))
        # the reference on the same face tensors: fp64 backward with fp64 weights (low noise)
        # and its own fp32 output, both scattered to the vertices in double
        r_out, r_mask, r_idx, ctxs = orc.dibr_rasterization_forward(H, W, z, img, feat, nz,
                                                                    ops=ref, **kw)
        assert torch.equal(r_idx, i_m) and torch.equal(r_out, o_m)
        ctx64 = orc.ctxs_to_double(ctxs)
        ctx64[0]["weights"] = pu.exact_weights64(
            H, W, img, r_idx, 1000. if kw.get("multiplier") is None else kw["multiplier"],
            1e-8 if kw.get("eps") is None else kw["eps"], saved_weights=ctxs[0]["weights"])
        gi64, gf64 = orc.dibr_rasterization_backward(ctx64, g1.double(), g2.double(), ops=ref)
        gi32, gf32 = orc.dibr_rasterization_backward(ctxs, g1, g2, ops=ref)
        vi64, vi32 = scatter64(gi64, faces, V), scatter64(gi32, faces, V)
        vf64, vf32 = scatter64(gf64, faces, V), scatter64(gf32, faces, V)
        if shared:
            vf64, vf32 = vf64.sum(0, keepdim=True), vf32.sum(0, keepdim=True)
        stats["grad_vertices_image"] = vertex_check("grad_vertices_image", vimg_l.grad, vi32, vi64)
        stats["grad_vertex_features"] = vertex_check("grad_vertex_features", vfeat_l.grad, vf32,
                                                     vf64)
    print("PARITY-MESH", stats)
    pu.log(stats)
    if grads:
        assert stats["grad_vertices_image"]["ok"] and stats["grad_vertex_features"]["ok"], stats
    return stats


def fixture_inputs(batch_size=3):
    g = wl.fixture_geometry(batch_size=batch_size)
    vcam = g["vertices_camera"].to(DEV).contiguous()
    proj = wl.generate_perspective_projection(3.14159265358979 / 4.).to(DEV)
    vimg = wl.perspective_camera(vcam, proj).contiguous()
    V = vcam.shape[1]
    vfeat = torch.rand((batch_size, V, 4), device=DEV, generator=torch.Generator(device=DEV).manual_seed(7))
    return dict(vertices_camera=vcam, vertices_image=vimg, faces=g["faces"].to(DEV),
                vertex_features=vfeat)


@pytest.mark.parametrize("H,W", [(64, 64), (37, 53), (200, 333)])
def test_fixture_mesh(kal, ref, H, W):
    s = compare(kal, ref, fixture_inputs(), H, W, "model.obj")
    assert s["covered"] > 100


def test_shared_features_and_parameters(kal, ref):
    compare(kal, ref, fixture_inputs(), 96, 80, "model.obj shared", shared=True, knum=5,
            boxlen=0.05, sigmainv=3000)


def test_cfg2_mesh(kal, ref):
    inp = wl.make_config("cfg2_70k_8x512", device=DEV)
    s = compare(kal, ref, inp, 512, 512, "cfg2 mesh", shared=True)
    assert s["covered"] > 100000 and s["soft"] > 1000


def test_cfg5_mesh_distinct_meshes(kal, ref):
    inp = wl.make_config("cfg5_256x2k_128", device=DEV, views=64)
    s = compare(kal, ref, inp, 128, 128, "cfg5 mesh")
    assert s["covered"] > 10000


def test_cfg3_mesh_two_views_forward(kal, ref):
    inp = wl.make_config("cfg3_200k_64x1024", device=DEV, views=64, view_slice=(8, 10))
    compare(kal, ref, inp, 1024, 1024, "cfg3 mesh[8:10]", shared=True)


def test_valid_faces_override_and_bad_ids(kal):
    inp = fixture_inputs(2)
    vcam, vimg, faces, vfeat = (inp[k] for k in ("vertices_camera", "vertices_image", "faces",
                                                 "vertex_features"))
    B, F = 2, faces.shape[0]
    valid = torch.rand((B, F), device=DEV) < 0.6
    z, img, feat, nz = face_tensors(kal, inp)
    (o_f, i_f) = kal.render.mesh.rasterize(48, 48, z, img, feat, valid_faces=valid)
    m_f = kal.render.mesh.dibr_soft_mask(img, i_f)
    o_m, m_m, i_m = kal.render.mesh.dibr_rasterization_mesh(48, 48, vcam, vimg, faces, vfeat,
                                                            valid_faces=valid)
    assert torch.equal(i_m, i_f) and torch.equal(o_m, o_f) and torch.equal(m_m, m_f)
    # faces with a vertex id outside [0, V) are skipped (int32 and int64 face lists alike)
    bad = faces.clone()
    bad[::7, 1] = vcam.shape[1] + 5
    bad[3::11, 0] = -1
    keep = ((bad >= 0) & (bad < vcam.shape[1])).all(dim=1)
    o_a, m_a, i_a = kal.render.mesh.dibr_rasterization_mesh(48, 48, vcam, vimg, bad.int(), vfeat)
    vmask = keep[None].expand(B, -1) & (nz >= 0)
    o_b, i_b = kal.render.mesh.rasterize(48, 48, z, img, feat, valid_faces=vmask)
    assert torch.equal(i_a, i_b) and torch.equal(o_a, o_b)


def test_graph_replay_matches_eager(kal):
    inp = wl.make_config("cfg5_256x2k_128", device=DEV, views=8)
    vcam, vimg, faces, vfeat = (inp[k] for k in ("vertices_camera", "vertices_image", "faces",
                                                 "vertex_features"))
    vfeat = vfeat.contiguous()
    faces32 = faces.int()
    g1 = torch.rand((8, 128, 128, 3), device=DEV)
    g2 = torch.rand((8, 128, 128), device=DEV)

    def run(graph):
        a = vimg.detach().requires_grad_(True)
        f = vfeat.detach().requires_grad_(True)
        o, m, i = kal.render.mesh.dibr_rasterization_mesh(128, 128, vcam, a, faces32, f,
                                                          cuda_graph=graph)
        torch.autograd.backward((o, m), (g1, g2))
        return o.clone(), m.clone(), i.clone(), a.grad.clone(), f.grad.clone()

    want = run(False)
    for _ in range(4):                 # first sighting eager, then captured, then replayed
        got = run(True)
        for x, y in zip(got[:3], want[:3]):
            assert torch.equal(x, y)
        for x, y in zip(got[3:], want[3:]):
            assert float((x - y).abs().max()) <= 1e-5 * float(y.abs().max())


# ------------------------------------------------- vertices + cameras in (one call)
def camera_inputs(cfg="cfg5_256x2k_128", views=8, shared=False):
    inp = wl.make_config(cfg, device=DEV, views=views)
    rot, trans, proj = inp["camera"]
    verts = inp["vertices"].contiguous()
    vfeat = inp["vertex_features"].contiguous()
    if shared:
        verts, vfeat = verts[:1].contiguous(), vfeat[:1].contiguous()
    return inp, verts, vfeat, rot.contiguous(), trans.contiguous(), proj.contiguous()


@pytest.mark.parametrize("shared", [False, True])
@pytest.mark.parametrize("mode", ["rot_trans", "transform"])
def test_vertex_transform_matches_torch(kal, shared, mode):
    """camera transform + projection kernel against the torch restatement of
    rotate_translate_points / perspective_camera (camera/legacy.py:22-37,120-139), forward
    and gradient."""
    from kaolin_windows_b200.render.mesh.dibr import VertexTransformB200
    _, verts, _, rot, trans, proj = camera_inputs(shared=shared)
    B = rot.shape[0]
    tf = None
    if mode == "transform":
        tf = torch.cat([rot.permute(0, 2, 1), (-trans.unsqueeze(1) @ rot.permute(0, 2, 1))], dim=1)
        tf = tf.contiguous()
    v1 = verts.clone().requires_grad_(True)
    if tf is None:
        vcam, vimg = VertexTransformB200.apply(v1, proj, rot, trans, None, B)
    else:
        vcam, vimg = VertexTransformB200.apply(v1, proj, None, None, tf, B)
    v2 = verts.clone().double().requires_grad_(True)
    ve = v2.expand(B, -1, -1)
    if tf is None:
        rcam = wl.rotate_translate_points(ve, rot.double(), trans.double())
    else:
        rcam = torch.nn.functional.pad(ve, (0, 1), value=1.) @ tf.double()
    rimg = wl.perspective_camera(rcam, proj.double())
    assert torch.allclose(vcam.double(), rcam, rtol=1e-5, atol=1e-5)
    assert torch.allclose(vimg.double(), rimg, rtol=1e-5, atol=1e-5)
    gc, gi = torch.rand_like(vcam), torch.rand_like(vimg)
    torch.autograd.backward((vcam, vimg), (gc, gi))
    torch.autograd.backward((rcam, rimg), (gc.double(), gi.double()))
    assert v1.grad.shape == verts.shape
    assert float((v1.grad.double() - v2.grad).abs().max()) <= 1e-5 * float(v2.grad.abs().max())


@pytest.mark.parametrize("shared", [False, True])
def test_vertices_call_equals_mesh_call(kal, shared):
    """dibr_rasterization_vertices == VertexTransform + dibr_rasterization_mesh (bit for bit),
    eager and replayed from CUDA graphs; gradients reach vertices and vertex features."""
    from kaolin_windows_b200.render.mesh.dibr import VertexTransformB200
    inp, verts, vfeat, rot, trans, proj = camera_inputs(shared=shared)
    faces = inp["faces"].int()
    B = rot.shape[0]
    g1 = torch.rand((B, 128, 128, 3), device=DEV)
    g2 = torch.rand((B, 128, 128), device=DEV)
    v1 = verts.clone().requires_grad_(True)
    f1 = vfeat.clone().requires_grad_(True)
    vcam, vimg = VertexTransformB200.apply(v1, proj, rot, trans, None, B)
    o_m, m_m, i_m = kal.render.mesh.dibr_rasterization_mesh(128, 128, vcam, vimg, faces, f1)
    torch.autograd.backward((o_m, m_m), (g1, g2))
    # the torch camera chain lands within rounding of the kernel's: same image up to a few
    # boundary pixels
    t_cam = wl.rotate_translate_points(verts.expand(B, -1, -1), rot, trans)
    o_t, m_t, i_t = kal.render.mesh.dibr_rasterization_mesh(
        128, 128, t_cam.contiguous(), wl.perspective_camera(t_cam, proj).contiguous(), faces,
        vfeat.expand(B, -1, -1).contiguous())
    assert float((i_t != i_m).float().mean()) < 1e-4
    for graph in (False, True, True, True):
        v2 = verts.clone().requires_grad_(True) if not graph else v_static
        f2 = vfeat.clone().requires_grad_(True) if not graph else f_static
        if graph:
            v2.grad = f2.grad = None
        o, m, i = kal.render.mesh.dibr_rasterization_vertices(
            128, 128, v2, faces, f2, proj, rot, trans, cuda_graph=graph)
        torch.autograd.backward((o, m), (g1, g2))
        assert torch.equal(i, i_m) and torch.equal(o, o_m) and torch.equal(m, m_m)
        for a, b in ((v2.grad, v1.grad), (f2.grad, f1.grad)):
            assert a.shape == b.shape
            assert float((a - b).abs().max()) <= 1e-5 * float(b.abs().max())
        if not graph:
            v_static = verts.clone().requires_grad_(True)
            f_static = vfeat.clone().requires_grad_(True)
