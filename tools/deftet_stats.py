#!/usr/bin/env python
"""Diagnostic: the cell grid deftet_sparse_render builds for the two workloads of
tools/bench_deftet.py (grid size, wide-list length, candidates tested per sample point).
Reads the library's workspace after a forward call (layout: csrc/dibr_deftet.cu,
deftet_layout)."""
import ctypes
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import kaolin_windows_b200 as kal      # noqa: E402
import workloads as wl                 # noqa: E402
from test_deftet_gpu import random_soup  # noqa: E402

dev = torch.device("cuda", 0)
lib = kal._lib.lib()
KW = 16


def up256(v):
    return (v + 255) // 256 * 256


def run(name, z, img, feat, P, K):
    B, F = z.shape[:2]
    D = feat.shape[-1]
    g = torch.Generator(dev).manual_seed(0)
    pix = torch.rand((B, P, 2), device=dev, generator=g) * 2 - 1
    rng = torch.tensor([[[-10., 0.]]], device=dev).repeat(B, P, 1)
    n = lib.dibr_deftet_workspace_bytes(B, F, P)
    ws = torch.zeros(n, dtype=torch.uint8, device=dev)
    idx = torch.empty((B, P, K), dtype=torch.long, device=dev)
    w = torch.empty((B, P, K, 3), device=dev)
    out = torch.empty((B, P, K, D), device=dev)
    p = lambda t: ctypes.c_void_p(t.data_ptr())
    rc = lib.dibr_deftet_fwd(0, B, F, P, K, D, p(z), p(img), p(feat), p(pix), p(rng), 1e-8,
                             p(idx), p(out), p(w), p(ws), n, None)
    assert rc == 0
    torch.cuda.synchronize()
    cap = min(512 * 512, 4 * P + 16)
    raw = ws.cpu().numpy()
    SB = 160                                       # sizeof(ViewStats)
    stats = raw[:B * SB].view(np.uint32).reshape(B, SB // 4)
    off_count = up256(B * SB)
    counts = raw[off_count:off_count + B * cap * 4].view(np.int32).reshape(B, cap)
    off_grids = up256(off_count + B * cap * 4)
    grids = raw[off_grids:off_grids + B * 32].view(np.int32).reshape(B, 8)
    res = {"cells_cap": cap}
    for b in range(min(B, 2)):
        G = int(grids[b, 6])
        c = counts[b][:G * G]
        nz = c[c > 0]
        ent = stats[b, 8:28].copy().view(np.uint64)
        res["view%d" % b] = {
            "G": G, "wide": int(stats[b, 4]),
            "entries_per_level": [int(e) for e in ent],
            "wides_per_level": [int(w) for w in stats[b, 28:38]],
            "cells_used": int((c > 0).sum()), "entries": int(c.sum()),
            "expected_list_len_uniform_sample": float(c.sum()) / (G * G),
            "list_len_mean_nonempty": float(nz.mean()) if len(nz) else 0,
            "list_len_p99": float(np.percentile(nz, 99)) if len(nz) else 0,
            "list_len_max": int(c.max())}
    return name, res


res = {}
inp = wl.make_config("cfg2_70k_8x512", device=dev)
n, r = run("cfg2", inp["face_vertices_z"], inp["face_vertices_image"], inp["face_features"],
           16384, 30)
res[n] = r
z, img, feat = random_soup(100000, 2, torch.float32, seed=0, size=0.15)
n, r = run("soup", z, img, feat[..., :3].contiguous(), 16384, 100)
res[n] = r
print(json.dumps(res, indent=1))
