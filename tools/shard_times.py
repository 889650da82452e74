#!/usr/bin/env python
"""Device time of one fwd+bwd step (CUDA graphs, L2 flushed) for each rank's shard of the
bench's N-GPU job, all on one GPU: how unequal the shards of the weak-scaling run are
(the N-GPU step time is the max over ranks plus the gradient exchange)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import kaolin_windows_b200 as kal      # noqa: E402
import workloads as wl                 # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 8
dev = torch.device("cuda", 0)
CFG, VPG = "cfg2_70k_8x512", 8
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
out = {}
for rank in range(N):
    lo, hi = kal.parallel.shard_range(VPG * N, rank, N)
    inp = wl.make_config(CFG, device=dev, views=VPG * N, view_slice=(lo, hi))
    z, img, feat, nz = (inp[k] for k in ("face_vertices_z", "face_vertices_image",
                                         "face_features", "face_normals_z"))
    H, W = inp["H"], inp["W"]
    gen = torch.Generator(device=dev).manual_seed(1)
    g1 = torch.rand((VPG, H, W, 3), device=dev, generator=gen)
    g2 = torch.rand((VPG, H, W), device=dev, generator=gen)

    def step():
        i, f = img.detach().requires_grad_(True), feat.detach().requires_grad_(True)
        o, m, _ = kal.render.mesh.dibr_rasterization(H, W, z, i, f, nz, cuda_graph=True)
        torch.autograd.backward((o, m), (g1, g2))

    for _ in range(6):
        step()
    ts = []
    for _ in range(20):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); step(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    out["rank%d" % rank] = round(sum(ts) / len(ts), 4)
print(json.dumps({"n_ranks": N, "ms_per_step_by_shard": out}))
