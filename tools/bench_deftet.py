#!/usr/bin/env python
"""deftet_sparse_render timed beside the reference's own CUDA kernels
(oracle/_ref) wrapped in the reference's torch-side code, forward and forward + backward,
CUDA events, one GPU.  Two workloads: the config-2 mesh (69 938 faces, 8 views, two layers
per point) and a volumetric soup (100 000 triangles, tens of layers per point).
Prints one JSON object."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import kaolin_windows_b200 as kal      # noqa: E402
import workloads as wl                 # noqa: E402
from oracle import dibr_oracle as orc  # noqa: E402
from test_deftet_gpu import ref_forward, random_soup  # noqa: E402

dev = torch.device("cuda", 0)
ref = orc.load_ref_ops()


def timed(fn, n=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3     # us


def case(name, z, img, feat, P, knum, lo, hi):
    B = z.shape[0]
    g = torch.Generator(dev).manual_seed(0)
    pix = torch.rand((B, P, 2), device=dev, generator=g) * 2 - 1
    rng = torch.tensor([[[lo, hi]]], device=dev).repeat(B, P, 1)
    cot = torch.rand((B, P, knum, feat.shape[-1]), device=dev, generator=g)

    def mine(bw):
        i, f = img.detach().requires_grad_(bw), feat.detach().requires_grad_(bw)
        out, idx = kal.render.mesh.deftet_sparse_render(pix, rng, z, i, f, knum)
        if bw:
            out.backward(cot)
        return idx

    def theirs(bw):
        out, idx, w = ref_forward(ref, pix, rng, z, img, feat, knum)
        if bw:
            ref.deftet_sparse_render_backward_cuda(cot, idx, w, img, feat, 1e-8)
        return idx

    idx = mine(False)
    assert torch.equal(idx, theirs(False))
    hits = (idx >= 0).sum(-1)
    r = {"views": B, "faces": z.shape[1], "points_per_view": P, "knum": knum,
         "hits_per_point_mean": round(float(hits.float().mean()), 2),
         "hits_per_point_max": int(hits.max())}
    for bw in (False, True):
        k = "fwd+bwd" if bw else "fwd"
        r[k + " (us)"] = {"b200": round(timed(lambda: mine(bw)), 1),
                          "reference_kernels": round(timed(lambda: theirs(bw), n=3, warm=1), 1)}
    r["Mpoints/s fwd+bwd"] = round(B * P / r["fwd+bwd (us)"]["b200"], 1)
    return name, r


res = {}
ONLY = sys.argv[1] if len(sys.argv) > 1 else ""      # "cfg2" | "soup" | "" (both)
inp = wl.make_config("cfg2_70k_8x512", device=dev) if ONLY != "soup" else None
if inp is not None:
    n, r = case("cfg2 mesh, 16384 points/view, knum 30", inp["face_vertices_z"],
                inp["face_vertices_image"], inp["face_features"], 16384, 30, -10., 0.)
    res[n] = r
if ONLY != "cfg2":
    z, img, feat = random_soup(100000, 2, torch.float32, seed=0, size=0.15)
    n, r = case("soup 100k faces, 16384 points/view, knum 100", z, img,
                feat[..., :3].contiguous(), 16384, 100, -10., 0.)
    res[n] = r
print(json.dumps({"op": "deftet_sparse_render", "results": res}))
