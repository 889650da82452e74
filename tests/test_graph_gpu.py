"""dibr_rasterization(cuda_graph=True): the CUDA-graph replay must return exactly what the
eager call returns (same kernels, same buffers' layout) -- on first sighting (eager), on
capture, on replay, after the inputs changed in place, and after a workspace overflow."""
import pytest
import torch

import kaolin_windows_b200 as kal
from kaolin_windows_b200 import _workspace
from kaolin_windows_b200.render.mesh import dibr_graph
import workloads as wl

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _inputs(views=2, n=24, seed=0):
    v, f = wl.grid_sphere(n, n, noise=0.02, seed=seed)
    inp = wl.render_inputs(v, f, views, device=DEV)
    return [inp[k].contiguous() for k in ("face_vertices_z", "face_vertices_image",
                                          "face_features", "face_normals_z")]


def _run(z, img, feat, nz, H, W, g1, g2, graph):
    a = img.detach().requires_grad_(True)
    b = feat.detach().requires_grad_(True)
    out, mask, idx = kal.render.mesh.dibr_rasterization(H, W, z, a, b, nz, cuda_graph=graph)
    torch.autograd.backward((out, mask), (g1, g2))
    return out.clone(), mask.clone(), idx.clone(), a.grad.clone(), b.grad.clone()


def _same(got, want):
    assert torch.equal(got[2], want[2]) and torch.equal(got[0], want[0])
    assert torch.equal(got[1], want[1])
    for g, w in zip(got[3:], want[3:]):
        assert float((g - w).abs().max()) <= 1e-5 * float(w.abs().max()) + 1e-12


def test_graph_replay_matches_eager():
    dibr_graph.clear()
    H, W = 96, 80
    z, img, feat, nz = _inputs()
    gen = torch.Generator(device=DEV).manual_seed(2)
    g1 = torch.rand((z.shape[0], H, W, 3), device=DEV, generator=gen)
    g2 = torch.rand((z.shape[0], H, W), device=DEV, generator=gen)
    want = _run(z, img, feat, nz, H, W, g1, g2, False)
    for i in range(4):                       # eager, capture, replay, replay
        _same(_run(z, img, feat, nz, H, W, g1, g2, True), want)
    assert len(dibr_graph._entries) == 1
    e = next(iter(dibr_graph._entries.values()))
    assert e.graph is not None and len(e.bwd_graphs) == 1
    # new data at the same addresses: the replay must see it
    img.mul_(0.93)
    feat.add_(0.1)
    want2 = _run(z, img, feat, nz, H, W, g1, g2, False)
    assert not torch.equal(want2[1], want[1])
    _same(_run(z, img, feat, nz, H, W, g1, g2, True), want2)
    # only one cotangent, and cotangents at new addresses
    a = img.detach().requires_grad_(True)
    out, mask, idx = kal.render.mesh.dibr_rasterization(H, W, z, a, feat, nz, cuda_graph=True)
    mask.backward(g2.clone())
    a2 = img.detach().requires_grad_(True)
    kal.render.mesh.dibr_rasterization(H, W, z, a2, feat, nz)[1].backward(g2)
    assert float((a.grad - a2.grad).abs().max()) <= 1e-5 * float(a2.grad.abs().max())
    dibr_graph.clear()


def test_graph_stale_backward_raises():
    dibr_graph.clear()
    H, W = 64, 64
    z, img, feat, nz = _inputs(views=1, n=12)
    outs = []
    for _ in range(3):
        a = img.detach().requires_grad_(True)
        outs.append(kal.render.mesh.dibr_rasterization(H, W, z, a, feat, nz, cuda_graph=True))
    with pytest.raises(RuntimeError, match="overwritten"):
        outs[1][1].sum().backward()
    outs[2][1].sum().backward()
    dibr_graph.clear()


def test_graph_grows_workspace_on_overflow(monkeypatch):
    dibr_graph.clear()
    monkeypatch.setattr(_workspace, "initial_caps", lambda B, F, H=0, W=0, knum=0: [64, 64, 8])
    H, W = 96, 96
    z, img, feat, nz = _inputs(views=2, n=20, seed=3)
    gen = torch.Generator(device=DEV).manual_seed(5)
    g1 = torch.rand((2, H, W, 3), device=DEV, generator=gen)
    g2 = torch.rand((2, H, W), device=DEV, generator=gen)
    monkeypatch.undo()
    want = _run(z, img, feat, nz, H, W, g1, g2, False)
    monkeypatch.setattr(_workspace, "initial_caps", lambda B, F, H=0, W=0, knum=0: [64, 64, 8])
    for _ in range(5):
        _same(_run(z, img, feat, nz, H, W, g1, g2, True), want)
    e = next(iter(dibr_graph._entries.values()))
    assert e.caps[0] > 64 and e.caps[1] > 64 and e.caps[2] > 8 and e.graph is not None
    dibr_graph.clear()
