#!/usr/bin/env python
"""Distribution of the soft mask's (pixel, face) pairs over 8x4 regions for a BASELINE config
(reads the pair buffer's act_cnt table out of the workspace after one forward)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import kaolin_windows_b200 as kal      # noqa: E402
from kaolin_windows_b200 import _lib, _workspace   # noqa: E402
import workloads as wl                 # noqa: E402

cfg = sys.argv[1] if len(sys.argv) > 1 else "cfg2_70k_8x512"
views = int(sys.argv[2]) if len(sys.argv) > 2 else None
dev = torch.device("cuda", 0)
inp = wl.make_config(cfg, device=dev, views=views)
H, W = inp["H"], inp["W"]
z, img, feat, nz = (inp[k] for k in ("face_vertices_z", "face_vertices_image", "face_features",
                                     "face_normals_z"))
B, F = z.shape[:2]
got = {}
orig = _workspace.run


def spy(*a, **k):
    r = orig(*a, **k)
    got["r"] = r
    return r


_workspace.run = spy
import importlib
dibr = importlib.import_module("kaolin_windows_b200.render.mesh.dibr")
for _ in range(2):
    out, mask, idx = kal.render.mesh.dibr_rasterization(H, W, z, img, feat, nz, 7000., 0.02, 30)
torch.cuda.synchronize()
ws, cap_r, cap_s, cap_p = got["r"]
lib = _lib.lib()
start = lib.dibr_workspace_bytes_for(B, F, H, W, cap_r, cap_s, 0, 30)
off_cnt = start + (cap_p * 4 + 255) // 256 * 256
cnt = ws[off_cnt:off_cnt + cap_p * 64].view(torch.int16).reshape(cap_p, 32).cpu().numpy().astype(np.int64)
cnt[cnt < 0] = 0        # 0xffff = covered pixel
tot = cnt.sum(1)
mx = cnt.max(1)
act = tot > 0
print("cap_p", cap_p, "active regions", int(act.sum()), "pairs", int(tot.sum()),
      "uncovered px with pairs", int((cnt > 0).sum()))
print("pairs/region percentiles 10/50/90/99/max:", np.percentile(tot[act], [10, 50, 90, 99, 100]))
print("max hits per region 10/50/90/99/max:", np.percentile(mx[act], [10, 50, 90, 99, 100]))
h, e = np.histogram(tot[act], bins=[0, 32, 64, 128, 256, 512, 768, 961])
print("histogram of pairs/region", dict(zip(e[1:].tolist(), h.tolist())))
print("share of pairs in regions > 512:", tot[tot > 512].sum() / tot.sum())
