#!/usr/bin/env python
"""A few eager fwd+bwd steps of one BASELINE config on cuda:0 -- the short command ncu
profiles (launch list / --set full).

    one_step.py [cfg] [views] [steps] [mesh|faces]

mesh (default): bench.py's step, render.mesh.dibr_rasterization_vertices (vertices + cameras
in, vertex-level gradients out); faces: the reference-shaped dibr_rasterization on face tensors."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import kaolin_windows_b200 as kal
import workloads as wl

cfg = sys.argv[1] if len(sys.argv) > 1 else "cfg2_70k_8x512"
views = int(sys.argv[2]) if len(sys.argv) > 2 else None
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
api = sys.argv[4] if len(sys.argv) > 4 else "mesh"
inp = wl.make_config(cfg, device="cuda:0", views=views)
H, W = inp["H"], inp["W"]
z, img, feat, nz = (inp[k] for k in ("face_vertices_z", "face_vertices_image", "face_features",
                                     "face_normals_z"))
rot, trans, proj = (t.contiguous() for t in inp["camera"])
shared = wl.CONFIGS[cfg]["mesh"][0] != "sphere_batch"
verts, vfeat = inp["vertices"].contiguous(), inp["vertex_features"].contiguous()
if shared:
    verts, vfeat = verts[:1].contiguous(), vfeat[:1].contiguous()
faces32 = inp["faces"].int().contiguous()
gen = torch.Generator(device="cuda:0").manual_seed(1234)
g1 = torch.rand((z.shape[0], H, W, feat.shape[-1]), device="cuda:0", generator=gen)
g2 = torch.rand((z.shape[0], H, W), device="cuda:0", generator=gen)
for _ in range(steps):
    if api == "mesh":
        v = verts.detach().requires_grad_(True)
        f = vfeat.detach().requires_grad_(True)
        out, mask, idx = kal.render.mesh.dibr_rasterization_vertices(
            H, W, v, faces32, f, proj, rot, trans, sigmainv=7000, boxlen=0.02, knum=30)
    else:
        a = img.detach().requires_grad_(True)
        f = feat.detach().requires_grad_(True)
        out, mask, idx = kal.render.mesh.dibr_rasterization(H, W, z, a, f, nz, 7000, 0.02, 30)
    torch.autograd.backward((out, mask), (g1, g2))
torch.cuda.synchronize()
print("ok", cfg, api, tuple(z.shape), int((idx >= 0).sum()))
