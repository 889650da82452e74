#!/usr/bin/env python
"""Development aid: per-CTA start/end/SM trace of one instrumented kernel on cfg2.

Build with  NVCC_EXTRA=-DDIBR_TRACE sh kaolin-windows_b200/csrc/build.sh  first; the
kernel named on the command line must contain DIBR_TRACE_BEGIN/END.  Usage:
    python tools/trace_ctas.py fwd|bwd
(all instrumented kernels of that pass write the same buffer: instrument one at a time).
"""
import ctypes
import sys

import numpy as np
import torch

sys.path.insert(0, __file__.rsplit("/", 2)[0])
import workloads as wl                      # noqa: E402
import kaolin_windows_b200 as kal           # noqa: E402
from kaolin_windows_b200 import _lib        # noqa: E402

which = sys.argv[1] if len(sys.argv) > 1 else "fwd"
dev = torch.device("cuda:0")
inp = wl.make_config("cfg2_70k_8x512", device=dev)
z, img, feat, nz = (inp[k] for k in ("face_vertices_z", "face_vertices_image", "face_features",
                                     "face_normals_z"))
H = W = 512
lib = ctypes.CDLL(_lib.LIB_PATH)
lib.dibr_trace_set.argtypes = [ctypes.c_void_p]
NCTA = 1 << 16
buf = torch.zeros(NCTA * 4, dtype=torch.int64, device=dev)


def run(trace):
    img_l = img.clone().requires_grad_(True)
    feat_l = feat.clone().requires_grad_(True)
    if trace and which == "fwd":
        lib.dibr_trace_set(buf.data_ptr())
    out, mask, idx = kal.render.mesh.dibr_rasterization(H, W, z, img_l, feat_l, nz)
    lib.dibr_trace_set(None)
    loss = out.sum() + mask.sum()
    if trace and which == "bwd":
        lib.dibr_trace_set(buf.data_ptr())
    loss.backward()
    torch.cuda.synchronize()
    lib.dibr_trace_set(None)


for _ in range(3):
    run(False)
run(True)
d = buf.cpu().numpy().reshape(-1, 4)
d = d[d[:, 1] > 0]
t0 = d[:, 0].min()
dur = (d[:, 1] - d[:, 0]) / 1e3
print("CTAs traced %d; kernel span %.1f us; CTA dur mean %.2f max %.1f sum %.0f us" % (
    len(d), (d[:, 1].max() - t0) / 1e3, dur.mean(), dur.max(), dur.sum()))
print("duration histogram (us):", np.histogram(dur, bins=[0, 2, 5, 10, 20, 40, 80, 160, 1000]))
order = np.argsort(-dur)[:12]
for i in order:
    print("  dur %.1f us, start %.1f us, sm %d, user value %d" % (dur[i], (d[i, 0] - t0) / 1e3, d[i, 2], d[i, 3]))
if len(sys.argv) > 2 and sys.argv[2] == "phases":   # -DDIBR_TRACE_PHASES build of raster_forward
    u = d[:, 3].astype(np.uint64)
    ph = np.stack([(u >> np.uint64(42)) & np.uint64(16383), (u >> np.uint64(28)) & np.uint64(16383),
                   (u >> np.uint64(14)) & np.uint64(16383), u & np.uint64(16383)], 1).astype(np.float64) * 64 / 1.965e3
    tot = dur.sum()
    heavy = dur > 10
    print("phases, CTAs over 10 us only (%d): mean us per CTA %s of %.1f" % (
        heavy.sum(), np.round(ph[heavy].mean(0), 2), dur[heavy].mean()))
    print("phase time summed over CTAs (us): A %.0f  A2 %.0f  B %.0f  resolve(+tail) %.0f  | all %.0f" % (
        ph[:, 0].sum(), ph[:, 1].sum(), ph[:, 2].sum(), ph[:, 3].sum(), tot))
    print("  shares: A %.2f A2 %.2f B %.2f resolve %.2f other %.2f" % tuple(
        list(ph.sum(0) / tot) + [1 - ph.sum() / tot]))
span = (d[:, 1].max() - t0) / 1e3
busy = np.zeros(int(d[:, 2].max()) + 1)
np.add.at(busy, d[:, 2], dur)
print("per-SM summed CTA time / span: mean %.2f max %.2f (resident CTAs per SM on average)" % (
    (busy / span).mean(), (busy / span).max()))
# time by which X% of the CTA work is done
end = np.sort((d[:, 1] - t0) / 1e3)
print("CTA end-time percentiles 50/90/99/100: ", np.percentile(end, [50, 90, 99, 100]))
# when CTAs start, and how long the ones started in each window live
st = (d[:, 0] - t0) / 1e3
edges = np.arange(0, span + 5, 5)
for lo, hi in zip(edges[:-1], edges[1:]):
    m = (st >= lo) & (st < hi)
    if m.any():
        print("  started in [%2.0f, %2.0f) us: %5d CTAs, mean dur %.2f us, of them > 5 us: %d" % (
            lo, hi, m.sum(), dur[m].mean(), (dur[m] > 5).sum()))
