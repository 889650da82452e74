#!/usr/bin/env python
"""Where the host time of one fwd+bwd step goes (cProfile over N steps, GPU kept busy)."""
import cProfile
import os
import pstats
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import kaolin_windows_b200 as kal   # noqa: E402
import workloads as wl              # noqa: E402

cfg = sys.argv[1] if len(sys.argv) > 1 else "cfg2_70k_8x512"
views = int(sys.argv[2]) if len(sys.argv) > 2 else None
GRAPH = len(sys.argv) > 3 and sys.argv[3] == "graph"
dev = torch.device("cuda", 0)
inp = wl.make_config(cfg, device=dev, views=views)
H, W = inp["H"], inp["W"]
z, img, feat, nz = (inp[k] for k in ("face_vertices_z", "face_vertices_image", "face_features",
                                     "face_normals_z"))
B, D = z.shape[0], feat.shape[-1]
g1 = torch.rand((B, H, W, D), device=dev)
g2 = torch.rand((B, H, W), device=dev)


def step():
    a = img.detach().requires_grad_(True)
    b = feat.detach().requires_grad_(True)
    out, mask, idx = kal.render.mesh.dibr_rasterization(H, W, z, a, b, nz, 7000., 0.02, 30,
                                                        cuda_graph=GRAPH)
    torch.autograd.backward((out, mask), (g1, g2))


for _ in range(10):
    step()
torch.cuda.synchronize()
N = 300
t0 = time.perf_counter()
for _ in range(N):
    step()
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print("host loop %.1f us/step, with final sync %.1f us/step" % ((t1 - t0) / N * 1e6, (t2 - t0) / N * 1e6))
pr = cProfile.Profile()
pr.enable()
for _ in range(N):
    step()
pr.disable()
torch.cuda.synchronize()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(22)
