"""Scratch-workspace management for the C ABI's workspace protocol (include/dibr_b200.h).

The per-bin face lists have a data-dependent length.  Rather than synchronising to size
them exactly (the reference's wrapper synchronises too: torch.where at
rasterization.py:325), a call is issued optimistically with the capacity that was enough
last time for this problem shape; the library reports the needed entry counts into pinned
host memory right after its counting pass (an async 16-byte copy) and the wrapper reads
them while the GPU is already running the remaining kernels.  Only if a list overflowed is
the capacity grown and the call re-issued.
"""
import os
import threading
import time

import torch

from . import _lib

_caps = {}            # (device index, kind, B, F, H, W, knum) -> [raster_cap, soft_cap, pair_cap]
_tls = threading.local()
SYNC_MODE = os.environ.get("DIBR_B200_WORKSPACE", "poll")   # "poll" | "sync"
# soft-mask pair buffer: "1" (default) keeps compact per-pixel face lists for the backward,
# "0" always uses the self-contained (re-walking) kernels
USE_PAIRS = os.environ.get("DIBR_B200_PAIRS", "1") != "0"
PAIR_CAP_OVERRIDE = None   # tests: force a pair capacity (e.g. 1 to exercise the fallback)
PAIR_MAX_K = 1024          # kPairMaxK in csrc/dibr_common.cuh


def _pinned_slot(key):
    """int64[4] pinned report slot of one problem shape: {raster, soft} list entries needed
    (read in the same call) and {segments, regions} of the pair buffer (arrive later; read
    at the next call of the same shape)."""
    slots = getattr(_tls, "slots", None)
    if slots is None:
        slots = _tls.slots = {}
    t = slots.get(key)
    if t is None:
        t = slots[key] = torch.full((4,), -1, dtype=torch.int64).pin_memory()
    return t


def initial_caps(B, F, H=0, W=0, knum=0):
    n = max(B * F, 1)
    return [2 * n + 4096, 4 * n + 4096, B * H * W * knum // 2048 + 1024]


def run(kind, device, B, F, H, W, use_r, use_s, call, knum=0):
    """call(ws_ptr, ws_bytes, cap_r, cap_s, cap_p, needed_ptr) -> rc.  Returns
    (ws, cap_r, cap_s, cap_p) with ws the uint8 workspace tensor the call filled."""
    lib = _lib.lib()
    key = (device.index, kind, B, F, H, W, knum)
    caps = _caps.get(key)
    if caps is None:
        caps = _caps[key] = initial_caps(B, F, H, W, knum)
    use_p = USE_PAIRS and use_s and 0 < knum <= PAIR_MAX_K
    if use_p and SYNC_MODE != "sync":
        # what the previous call of this shape wanted (its report has long arrived)
        prev = _pinned_slot(key)
        want = max(int(prev[2]), int(prev[3]))
        if want > caps[2]:
            caps[2] = int(want * 1.25) + 1024
    for _attempt in range(4):
        cap_r = caps[0] if use_r else 0
        cap_s = caps[1] if use_s else 0
        cap_p = (caps[2] if PAIR_CAP_OVERRIDE is None else PAIR_CAP_OVERRIDE) if use_p else 0
        cap_p = min(cap_p, 1 << 23)
        nbytes = lib.dibr_workspace_bytes_for(B, F, H, W, cap_r, cap_s, cap_p, knum)
        if nbytes < 0:
            _lib.check(int(nbytes), "dibr_workspace_bytes_for")
        ws = torch.empty(int(nbytes), dtype=torch.uint8, device=device)
        if SYNC_MODE == "sync":
            rc = call(ws.data_ptr(), nbytes, cap_r, cap_s, cap_p, None)
            if rc == _lib.DIBR_OK:
                return ws, cap_r, cap_s, cap_p
            if rc != _lib.DIBR_E_WORKSPACE:
                _lib.check(rc, "libdibr_b200")
            # message carries the needed counts; fall through to generic growth
            caps[0] *= 2
            caps[1] *= 2
            continue
        slot = _pinned_slot(key)
        slot.fill_(-1)
        rc = call(ws.data_ptr(), nbytes, cap_r, cap_s, cap_p, slot.data_ptr())
        _lib.check(rc, "libdibr_b200")
        # the 16-byte copy lands shortly after the counting pass; the kernels behind it
        # keep the GPU busy while we look
        t0 = time.perf_counter()
        while True:
            nr, ns = int(slot[0]), int(slot[1])
            if nr >= 0 and ns >= 0:
                break
            if time.perf_counter() - t0 > 120.0:
                torch.cuda.synchronize(device)
                nr, ns = int(slot[0]), int(slot[1])
                if nr < 0 or ns < 0:
                    raise RuntimeError("libdibr_b200: list-size report never arrived")
                break
        if nr <= cap_r and ns <= cap_s:
            return ws, cap_r, cap_s, cap_p
        if use_r and nr > cap_r:
            caps[0] = int(nr * 1.25) + 4096
        if use_s and ns > cap_s:
            caps[1] = int(ns * 1.25) + 4096
    raise RuntimeError("libdibr_b200: workspace kept overflowing")
