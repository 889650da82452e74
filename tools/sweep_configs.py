#!/usr/bin/env python
"""Per-config fwd+bwd time and per-kernel breakdown on one GPU (development aid).

    python tools/sweep_configs.py [cfg:views ...]   e.g. cfg3_200k_64x1024:8 cfg4_1m_16x2048:2

Same step as bench.py (dibr_rasterization forward + backward with fixed cotangents),
CUDA-event timed, per-kernel times from the library's profiling hooks.
"""
import ctypes
import json
import sys

import torch

sys.path.insert(0, __file__.rsplit("/", 2)[0])
import workloads as wl                      # noqa: E402
import kaolin_windows_b200 as kal           # noqa: E402
from kaolin_windows_b200 import _lib        # noqa: E402
from bench import KERNELS, total_algorithmic_bytes  # noqa: E402

DEFAULT = ["cfg1_pizza_256:1", "cfg2_70k_8x512:8", "cfg3_200k_64x1024:8", "cfg4_1m_16x2048:2",
           "cfg5_256x2k_128:256"]
dev = torch.device("cuda:0")
lib = _lib.lib()
for spec in (sys.argv[1:] or DEFAULT):
    name, views = spec.split(":")
    inp = wl.make_config(name, device=dev, views=int(views))
    H, W = inp["H"], inp["W"]
    z, img, feat, nz = (inp[k] for k in ("face_vertices_z", "face_vertices_image",
                                         "face_features", "face_normals_z"))
    B, F, D = z.shape[0], z.shape[1], feat.shape[-1]
    g1 = torch.rand((B, H, W, D), device=dev)
    g2 = torch.rand((B, H, W), device=dev)

    def step(graph=True):
        img_l = img.detach().requires_grad_(True)
        feat_l = feat.detach().requires_grad_(True)
        out, mask, idx = kal.render.mesh.dibr_rasterization(H, W, z, img_l, feat_l, nz,
                                                            cuda_graph=graph)
        torch.autograd.backward((out, mask), (g1, g2))
        return idx, mask

    for _ in range(4):
        idx, mask = step()
    torch.cuda.synchronize()
    n = 10
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    for _ in range(3):            # eager calls of this shape first: capacities settle
        step(False)
    torch.cuda.synchronize()
    lib.dibr_profile_begin()
    for _ in range(3):
        step(False)
    cap = 256
    ids, t = (ctypes.c_int * cap)(), (ctypes.c_float * cap)()
    k = lib.dibr_profile_end(ids, t, cap)
    per = {}
    for i in range(k):
        per.setdefault(KERNELS[ids[i]], []).append(t[i])
    per = {a: round(sum(v) / len(v), 4) for a, v in per.items()}
    gb = total_algorithmic_bytes(B, F, H * W, D) / 1e9
    print(json.dumps({"config": name, "views": B, "faces": F, "HxW": [H, W], "ms_per_step": round(ms, 3),
                      "Mpix_s": round(B * H * W / ms / 1e3, 1),
                      "algorithmic_GB_s": round(gb / (ms * 1e-3), 1),
                      "covered": int((idx >= 0).sum()), "soft": int(((mask > 0) & (mask < 1)).sum()),
                      "kernel_ms": per}), flush=True)
    from kaolin_windows_b200.render.mesh import dibr_graph
    dibr_graph.clear()
    del inp, z, img, feat, nz, g1, g2
    torch.cuda.empty_cache()
