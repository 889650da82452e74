#!/usr/bin/env python
"""Run under torchrun (N >= 2 GPUs): the peer-memory exchange (csrc/dibr_peer.cu through
parallel.MeshGradExchange) against NCCL on the same gradients -- values, bit-identity across
ranks, repeated calls (eager, captured, replayed) -- and the time of both.
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/check_peer_exchange.py"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import kaolin_windows_b200 as kal      # noqa: E402
import workloads as wl                 # noqa: E402

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank)))
torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
v, f = wl.grid_sphere(187, 188, noise=0.02, seed=0)
faces = torch.as_tensor(f).long().to(dev)
V, F, B = v.shape[0], faces.shape[0], 8
gen = torch.Generator(device=dev).manual_seed(100 + rank)
gi = torch.randn((B, F, 3, 2), device=dev, generator=gen)
gf = torch.randn((B, F, 3, 3), device=dev, generator=gen)

os.environ["DIBR_B200_EXCHANGE"] = "nccl"
ex_nccl = kal.parallel.MeshGradExchange(faces, V)
ref = ex_nccl(gi, gf).clone()
os.environ["DIBR_B200_EXCHANGE"] = "peer"
ex_peer = kal.parallel.MeshGradExchange(faces, V)
ok = True
for it in range(6):                     # eager, eager + capture, replays
    gi.mul_(1.0 + 0.01 * it)
    want = ex_nccl(gi, gf).clone()
    got = ex_peer(gi, gf).clone()
    err = float((got - want).abs().max() / want.abs().max())
    gathered = [torch.empty_like(got) for _ in range(world)]
    dist.all_gather(gathered, got)
    same = all(torch.equal(gathered[0], g) for g in gathered)
    ok = ok and err < 1e-5 and same and ex_peer.mode == "peer"
    if rank == 0:
        print("iter", it, "mode", ex_peer.mode, "rel err vs nccl %.2e" % err,
              "bit-identical across ranks", same, flush=True)


def timed(ex, n=50):
    for _ in range(5):
        ex(gi, gf)
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        ex(gi, gf)
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / n * 1e3], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return round(float(t.item()), 1)


# ---- the exchange folded into the graphed backward of dibr_rasterization_vertices
# (exchange=): gradients of the shared mesh summed over ranks, against NCCL on this rank's
# own un-exchanged gradients; eager (first sighting), captured, replayed
lo, hi = kal.parallel.shard_range(4 * world, rank, world)
inp = wl.make_config("cfg5_256x2k_128", device=dev, views=4 * world, view_slice=(lo, hi))
rot, trans, proj = (t.contiguous() for t in inp["camera"])
verts = inp["vertices"][:1].contiguous()
vfeat = inp["vertex_features"][:1].contiguous()
faces32 = inp["faces"].int().contiguous()
g1 = torch.rand((hi - lo, 128, 128, 3), device=dev, generator=gen)
g2 = torch.rand((hi - lo, 128, 128), device=dev, generator=gen)
vex = kal.parallel.VertexGradExchange()


def render_step(exchange, graph):
    a = verts.detach().requires_grad_(True)
    b = vfeat.detach().requires_grad_(True)
    o, m, _ = kal.render.mesh.dibr_rasterization_vertices(128, 128, a, faces32, b, proj, rot, trans,
                                                          cuda_graph=graph, exchange=exchange)
    torch.autograd.backward((o, m), (g1, g2))
    return torch.cat([a.grad.reshape(-1), b.grad.reshape(-1)])


want = render_step(None, False).clone()
dist.all_reduce(want, op=dist.ReduceOp.SUM)
for it in range(5):
    got = render_step(vex, True)
    err = float((got - want).abs().max() / want.abs().max())
    gathered = [torch.empty_like(got) for _ in range(world)]
    dist.all_gather(gathered, got)
    same = all(torch.equal(gathered[0], g) for g in gathered)
    ok = ok and err < 1e-5 and same and vex.mode == "peer"
    if rank == 0:
        print("folded iter", it, "rel err vs nccl %.2e" % err, "bit-identical across ranks", same,
              flush=True)
got = render_step(vex, False)               # and the un-graphed form (SumGradsAcrossRanks)
ok = ok and float((got - want).abs().max() / want.abs().max()) < 1e-5
vex.release()

res = {"world": world, "ok": bool(ok), "exchange_us": {"nccl": timed(ex_nccl), "peer": timed(ex_peer)}}
if rank == 0:
    print(json.dumps(res), flush=True)
ex_nccl.release()
ex_peer.release()
torch.cuda.synchronize()
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
