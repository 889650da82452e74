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

# ---- SURVEY 8(f) rank 1 as a whole: vertices + cameras -> images -> vertex gradients.
# (a) the reference's chain with its torch code around our drop-in dibr_rasterization,
# (b) the same chain on our face-tensor kernels (round 1), (c) the mesh-level call (round 2)
H = W = 512
g1 = torch.rand((B, H, W, 3), device=dev)
g2 = torch.rand((B, H, W), device=dev)
verts1, colors1 = verts[:1].contiguous(), colors[:1].contiguous()
faces32 = faces.int().contiguous()


def chain(mod_prepare, mod_gather, graph):
    v = verts1.detach().requires_grad_(True)
    c = colors1.detach().requires_grad_(True)
    fvc, fvi, fn = mod_prepare(v.expand(B, -1, -1), faces, proj, camera_rot=rot, camera_trans=trans)
    ff = mod_gather(c.expand(B, -1, -1), faces)
    out, mask, _ = kal.render.mesh.dibr_rasterization(H, W, fvc[..., 2].contiguous(), fvi, ff,
                                                      fn[..., 2].contiguous(), cuda_graph=graph)
    torch.autograd.backward((out, mask), (g1, g2))


def mesh_level(graph):
    v = verts1.detach().requires_grad_(True)
    c = colors1.detach().requires_grad_(True)
    out, mask, _ = kal.render.mesh.dibr_rasterization_vertices(H, W, v, faces32, c, proj, rot, trans,
                                                               cuda_graph=graph)
    torch.autograd.backward((out, mask), (g1, g2))


res["vertices+cameras -> dibr_rasterization -> vertex gradients, fwd+bwd (us)"] = {
    "torch prepare_vertices/index + b200 dibr_rasterization (eager)":
        timed(lambda: chain(mo.prepare_vertices, mo.index_vertices_by_faces, False)),
    "b200 face-tensor kernels: prepare_vertices/index + dibr_rasterization (eager)":
        timed(lambda: chain(kal.render.mesh.prepare_vertices, kal.ops.mesh.index_vertices_by_faces, False)),
    "b200 mesh-level dibr_rasterization_vertices (eager)": timed(lambda: mesh_level(False)),
    "b200 mesh-level dibr_rasterization_vertices (cuda_graph)": timed(lambda: mesh_level(True))}

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
# ---- the same two consumer ops as bare C-ABI calls on preallocated buffers (device time:
# the public calls above are bound by Python / autograd-engine overhead per iteration)
import ctypes                           # noqa: E402
from kaolin_windows_b200 import _lib    # noqa: E402
lib = _lib.lib()
P = ctypes.c_void_p
st = P(torch.cuda.current_stream(dev).cuda_stream)
sums = torch.empty((B, 2), device=dev)
loss = torch.empty((), device=dev)
gl = torch.ones((), device=dev)
g_m1 = torch.empty_like(m1)


def iou_raw():
    lib.dibr_mask_iou_sums(B, H * W, P(m1.data_ptr()), P(m2.data_ptr()), P(sums.data_ptr()), st)
    lib.dibr_mask_iou_loss(B, P(sums.data_ptr()), P(loss.data_ptr()), st)
    lib.dibr_mask_iou_bwd(B, H * W, P(m1.data_ptr()), P(m2.data_ptr()), P(sums.data_ptr()),
                          P(gl.data_ptr()), P(g_m1.data_ptr()), None, st)


res["mask_iou fwd+bwd (us)"]["b200 kernels only (C ABI)"] = timed(iou_raw, 50)
out_t = torch.empty((B, H, W, 3), device=dev)
g_tex = torch.empty_like(tex)
g_uv = torch.empty_like(uv)
for mode_i, mode in enumerate(("nearest", "bilinear")):
    def tex_raw():
        lib.dibr_texture_mapping_fwd(B, H * W, 3, 512, 512, mode_i, P(uv.data_ptr()),
                                     P(tex.data_ptr()), P(out_t.data_ptr()), st)
        g_tex.zero_()
        lib.dibr_texture_mapping_bwd(B, H * W, 3, 512, 512, mode_i, P(uv.data_ptr()),
                                     P(tex.data_ptr()), P(cot_tex.data_ptr()), P(g_tex.data_ptr()),
                                     P(g_uv.data_ptr()) if mode_i else None, st)
    res["texture_mapping %s fwd+bwd (us)" % mode]["b200 kernels only (C ABI)"] = timed(tex_raw, 50)

res = {k: {a: round(b, 1) for a, b in v.items()} for k, v in res.items()}
print(json.dumps({"config": "cfg2_70k_8x512 sizes (8 views, 69938 faces, 512x512)", "results": res}))
