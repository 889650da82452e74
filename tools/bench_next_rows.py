#!/usr/bin/env python
"""Timings of the "next row" ops beside the reference's torch code
for the same step, forward + backward, at BASELINE config 2 sizes, CUDA events, one GPU.
The torch side is the restatement in oracle/mesh_oracle.py run on the GPU (the reference
functions are pure torch).  Prints one JSON object."""
import json
import math
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import kaolin_windows_b200 as kal      # noqa: E402
import workloads as wl                 # noqa: E402
from oracle import mesh_oracle as mo   # noqa: E402

dev = torch.device("cuda", 0)


def timed(fn, n=20):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3     # us


inp = wl.make_config("cfg2_70k_8x512", device=dev)
B = inp["face_vertices_z"].shape[0]
verts = inp["vertices"].expand(B, -1, -1).contiguous() if inp["vertices"].shape[0] == 1 \
    else inp["vertices"]
faces = inp["faces"]
V = verts.shape[1]
rot, trans, proj = (t.to(dev) for t in wl.circle_cameras(B))
colors = torch.rand((B, V, 3), device=dev)
res = {}


F = faces.shape[0]
cots = [torch.randn(s, device=dev) for s in ((B, F, 3, 3), (B, F, 3, 2), (B, F, 3), (B, F, 3, 3))]


def prep(mod_prepare, mod_gather, backward=True):
    v = verts.detach().requires_grad_(True)
    c = colors.detach().requires_grad_(True)
    fvc, fvi, fn = mod_prepare(v, faces, proj, camera_rot=rot, camera_trans=trans)
    ff = mod_gather(c, faces)
    if backward:
        torch.autograd.backward((fvc, fvi, fn, ff), cots)


for bw in (False, True):
    res["prepare_vertices+index_vertices_by_faces %s (us)" % ("fwd+bwd" if bw else "fwd")] = {
        "b200": timed(lambda: prep(kal.render.mesh.prepare_vertices,
                                   kal.ops.mesh.index_vertices_by_faces, bw)),
        "torch": timed(lambda: prep(mo.prepare_vertices, mo.index_vertices_by_faces, bw))}

H = W = 512
uv = torch.rand((B, H, W, 2), device=dev)
tex = torch.rand((B, 3, 512, 512), device=dev)
cot_tex = torch.randn((B, H, W, 3), device=dev)
for mode in ("nearest", "bilinear"):
    def tm(fn):
        u = uv.detach().requires_grad_(True)
        t = tex.detach().requires_grad_(True)
        fn(u, t, mode).backward(cot_tex)
    res["texture_mapping %s fwd+bwd (us)" % mode] = {
        "b200": timed(lambda: tm(kal.render.mesh.texture_mapping)),
        "torch": timed(lambda: tm(mo.texture_mapping))}

m1 = torch.rand((B, H, W), device=dev)
m2 = (torch.rand((B, H, W), device=dev) > 0.5).float()


def iou(fn):
    a = m1.detach().requires_grad_(True)
    fn(a, m2).backward()


res["mask_iou fwd+bwd (us)"] = {"b200": timed(lambda: iou(kal.metrics.render.mask_iou)),
                                "torch": timed(lambda: iou(mo.mask_iou))}
res = {k: {a: round(b, 1) for a, b in v.items()} for k, v in res.items()}
print(json.dumps({"config": "cfg2_70k_8x512 sizes (8 views, 69938 faces, 512x512)", "results": res}))
