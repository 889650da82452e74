"""Gradient parity bar shared by the GPU parity tests, and the PARITY log.

north_star asks for gradients "within 1e-4 relative" of the reference.  A tensor-wide
max-norm would let every small-magnitude face gradient be wrong by orders of magnitude, so
the bar is PER FACE (the reference's own tests compare elementwise:
tests/python/kaolin/render/mesh/test_rasterization.py:189-288):

    e_f = |g_f - g64_f|_inf / max(|g64_f|_inf, 1e-3 * |g64|_inf)        must be <= 1e-4

with g64 the reference's backward kernels run in fp64 for the discrete choices of the fp32
forward (selected faces, kept faces and nearest-feature types of the soft mask, and the fp32
soft-mask state, which is the backward op's input), and with the barycentric weights a
fp64 run of the reference's forward computes for those faces at those pixel centres
(`exact_weights64`, restating rasterization_cuda.cu:85-86,131-145 in double) -- i.e. "the
reference run in fp64"  without the pixels changing hands.  The fp32
weights the fp32 forward saves are NOT used for g64: the backward only needs them to
reconstruct the pixel centre (:302-303), and their rounding moves that centre by up to 1e-2
of a sub-pixel face -- that is the reference's own fp32 noise (`ref32_*` columns), not the
target.  The comparison against the fp64 backward on the saved fp32 weights (round 1's
reference) is logged beside it as `vs_saved_w`.

Where the reference's OWN fp32 result is further than 1e-4 from its fp64 result (1 / k3^2 amplification on sub-pixel faces, 1 - soft_mask cancellation),
a face may instead be no further from fp64 than 1.5x the reference's fp32 error -- capped
at 1e-3 -- and the number of faces that needed this clause is counted and logged.

Every comparison appends one JSON line to $DIBR_PARITY_LOG (default
gpurun_out/r2_parity.jsonl when that directory exists); the committed copy is
profiles/r2_parity.jsonl.
"""
import json
import os

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TOL = 1e-4          # north_star: gradients within 1e-4 relative
FLOOR = 1e-3        # faces whose gradient is below 1e-3 of the tensor's largest: absolute bar
ESCAPE_CAP = 1e-3   # hard cap of the "no worse than the reference's own fp32" clause


def _log_path():
    p = os.environ.get("DIBR_PARITY_LOG")
    if p:
        return p
    d = os.path.join(ROOT, "gpurun_out")
    return os.path.join(d, "r2_parity.jsonl") if os.path.isdir(d) else None


def log(record):
    p = _log_path()
    if p:
        with open(p, "a") as fh:
            fh.write(json.dumps(record) + "\n")


def per_face_err(g, g64):
    """(B, F) tensor of e_f for gradients shaped (B, F, 3, C)."""
    g, g64 = g.double(), g64.double()
    B, F = g.shape[:2]
    diff = (g - g64).abs().reshape(B, F, -1).amax(dim=2)
    mag = g64.abs().reshape(B, F, -1).amax(dim=2)
    floor = FLOOR * g64.abs().max().clamp_min(1e-300)
    return diff / torch.maximum(mag, floor)


def face_grad_check(name, mine, ref32, ref64):
    """Per-face bar.  Returns a stats dict; `ok` says whether every face met it."""
    e = per_face_err(mine, ref64)
    r = per_face_err(ref32, ref64)
    strict = e <= TOL
    escape = (~strict) & (e <= torch.clamp(1.5 * r, max=ESCAPE_CAP))
    bad = ~(strict | escape)
    touched = ref64.abs().reshape(e.shape[0], e.shape[1], -1).amax(dim=2) > 0
    n = e.numel()
    st = {
        "tensor": name, "faces": n, "faces_with_gradient": int(touched.sum()),
        "max_e": float(e.max()) if n else 0.0,
        "ref32_max_e": float(r.max()) if n else 0.0,
        "over_1e-4": int((~strict).sum()),
        "ref32_over_1e-4": int((r > TOL).sum()),
        "escape_used": int(escape.sum()), "failed": int(bad.sum()),
        "global_rel": float((mine.double() - ref64.double()).abs().max()
                            / ref64.double().abs().max().clamp_min(1e-300)) if n else 0.0,
    }
    if n and bool(bad.any()):
        w = int(torch.argmax(torch.where(bad, e, torch.zeros_like(e))))
        st["worst_face"] = [w // e.shape[1], w % e.shape[1]]
        st["worst_e"] = float(e.reshape(-1)[w])
        st["worst_ref32_e"] = float(r.reshape(-1)[w])
    st["ok"] = st["failed"] == 0
    return st


def exact_weights64(height, width, face_vertices_image, face_idx, multiplier, eps,
                    saved_weights=None):
    """(B, H, W, 3) float64: the weights the reference's forward kernel instantiated for
    double (rasterization_cuda.cu:85-86, 131-145) gives the face `face_idx` holds at each
    pixel; zeros for background.  The pixel centre is the FLOAT product the kernel forms
    (`multiplier / width * (2 * wididx + 1 - width)`, float x int), coordinates are
    double(fp32 input) * multiplier as RasterizeCuda.forward scales them (:337).
    Degenerate faces -- area term N = w0 + w1 + w2 below 1e-4 of |w0| + |w1| + |w2| (collinear
    corners: N is pure cancellation) or below 4 eps (the weights are w_k / +-eps, the sign
    decided by rounding) -- have no meaningful fp64 weights: there the fp32 forward's
    `saved_weights` (B, H, W, 3) are kept, as the reference's backward would see them."""
    dev = face_vertices_image.device
    B, H, W = face_idx.shape
    # fl32(multiplier / width) as an IEEE float DIVISION, formed with numpy: torch divides a
    # tensor by a host scalar as a multiplication by fl32(1 / width), which is 1 ulp off for
    # sizes that are not powers of two (520, 600: the centre moves by 3e-5 multiplied units,
    # 1e-4 of a sub-pixel face)
    xstep = torch.tensor(float(np.float32(multiplier) / np.float32(width)), dtype=torch.float32,
                         device=dev)
    ystep = torch.tensor(float(np.float32(multiplier) / np.float32(height)), dtype=torch.float32,
                         device=dev)
    xs = (xstep * torch.arange(1 - width, width + 1, 2, device=dev,
                               dtype=torch.int32).float()).double()
    ys = (ystep * torch.arange(height - 1, -height - 1, -2, device=dev,
                               dtype=torch.int32).float()).double()
    out = torch.zeros((B, H, W, 3), dtype=torch.float64, device=dev)
    b, y, x = torch.nonzero(face_idx >= 0, as_tuple=True)
    v = face_vertices_image.double()[b, face_idx[b, y, x]] * float(multiplier)   # (n, 3, 2)
    x0, y0 = xs[x], ys[y]
    ex = v[..., 0] - x0[:, None]
    ey = v[..., 1] - y0[:, None]
    w0 = ex[:, 1] * ey[:, 2] - ey[:, 1] * ex[:, 2]
    w1 = ex[:, 2] * ey[:, 0] - ey[:, 2] * ex[:, 0]
    w2 = ex[:, 0] * ey[:, 1] - ey[:, 0] * ex[:, 1]
    norm = w0 + w1 + w2
    healthy = (norm.abs() > 1e-4 * (w0.abs() + w1.abs() + w2.abs())) & (norm.abs() > 4.0 * float(eps))
    norm = norm + torch.copysign(torch.full_like(norm, float(eps)), norm)
    w = torch.stack([w0 / norm, w1 / norm, w2 / norm], dim=-1)
    if saved_weights is not None:
        w = torch.where(healthy[:, None], w, saved_weights.double()[b, y, x])
    out[b, y, x] = w
    return out
