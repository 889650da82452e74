"""world_size-2 gloo test of the multi-GPU plumbing (host-side logic only): view sharding
covers the batch exactly once, and the flat all-reduce of the shared mesh's gradients equals
the gradient of the un-sharded job.  The per-rank "render" is the CPU oracle here (this is
a test of the exchange step, not of the kernels)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import workloads as wl


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _vertex_grads(inp, H, W, lo, hi):
    """oracle fwd+bwd on views [lo, hi), face grads scattered to the shared vertices."""
    from oracle import dibr_oracle as orc
    z, img, feat, nz = (inp[k][lo:hi] for k in ("face_vertices_z", "face_vertices_image",
                                                "face_features", "face_normals_z"))
    out, mask, idx, ctxs = orc.dibr_rasterization_forward(H, W, z, img, feat, nz)
    g = torch.Generator().manual_seed(5)
    g1 = torch.rand((inp["face_vertices_z"].shape[0], H, W, feat.shape[-1]), generator=g)[lo:hi]
    g2 = torch.rand((inp["face_vertices_z"].shape[0], H, W), generator=g)[lo:hi]
    gi, gf = orc.dibr_rasterization_backward(ctxs, g1.contiguous(), g2.contiguous())
    V = inp["vertices"].shape[1]
    gv = torch.zeros(V, 2 + gf.shape[-1])
    gv.index_add_(0, inp["faces"].reshape(-1), torch.cat([gi, gf], dim=-1).sum(0).reshape(-1, gv.shape[1]))
    return gv


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from kaolin_windows_b200.parallel import shard_range, allreduce_mesh_grads
    v, f = wl.grid_sphere(12, 12, noise=0.02)
    inp = wl.render_inputs(v, f, 4)
    lo, hi = shard_range(4, rank, world)
    gv = _vertex_grads(inp, 48, 48, lo, hi)
    gv2 = gv[:, :2].clone().contiguous()
    gf2 = gv[:, 2:].clone().contiguous()
    allreduce_mesh_grads([gv2, gf2])
    if rank == 0:
        q.put(torch.cat([gv2, gf2], dim=1))
    dist.barrier()
    dist.destroy_process_group()


def test_view_sharded_allreduce_matches_unsharded():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    v, f = wl.grid_sphere(12, 12, noise=0.02)
    inp = wl.render_inputs(v, f, 4)
    want = _vertex_grads(inp, 48, 48, 0, 4)
    assert want.abs().max() > 0
    assert torch.allclose(got, want, rtol=1e-4, atol=1e-5 * float(want.abs().max()))
