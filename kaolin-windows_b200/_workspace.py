"""Scratch-workspace management for the C ABI's workspace protocol (include/dibr_b200.h).

The per-bin face lists have a data-dependent length.  Rather than synchronising to size
them exactly (the reference's wrapper synchronises too: torch.where at
rasterization.py:325), a call is issued optimistically with the capacity that was enough
last time for this problem shape; the library reports the needed entry counts into pinned
host memory right after its counting pass and the wrapper looks at them:

* ``DIBR_B200_WORKSPACE=lazy`` (default): NO host wait.  The first call of a shape is
  verified before it returns (and sizes the lists with 50 % headroom); after that a call
  returns as soon as it is queued, and its report is checked when a later call of the same
  shape (or ``check_pending()``) finds it arrived -- up to 8 calls may be in flight.  A list
  that grew past its capacity in the meantime is detected then: the capacity is raised and a
  RuntimeError names the earlier call whose result was incomplete (capacities are also
  raised ahead of need, at 80 % fill, so this takes a > 25 % jump between two calls).
* ``poll``: every call waits (spinning on the pinned slot) until its own report has arrived,
  i.e. until everything queued before it plus its binning has run; an overflow re-issues the
  call transparently.  Never wrong, but the host can never be more than one call ahead.
* ``sync``: the library synchronises the stream itself (needed_host = NULL).
"""
import os
import threading
import time

import torch

from . import _lib

_caps = {}            # (device index, kind, B, F, H, W, knum) -> [raster_cap, soft_cap, pair_cap]
_tls = threading.local()
SYNC_MODE = os.environ.get("DIBR_B200_WORKSPACE", "lazy")   # "lazy" | "poll" | "sync"
RING = 8              # lazy mode: calls in flight (unverified) per problem shape
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


class _Lazy:
    """Per-shape state of the lazy mode: a ring of pinned report slots and the calls whose
    reports have not been looked at yet."""

    def __init__(self):
        self.calibrated = False
        self.slots = [torch.full((4,), -1, dtype=torch.int64).pin_memory() for _ in range(RING)]
        self.next = 0
        self.pending = []          # [(slot, cap_r, cap_s)] in issue order

    def verify(self, caps, use_r, use_s, wait_all=False, block_oldest=False):
        """Looks at the reports that have arrived (in issue order).  Raises for a call that
        overflowed; raises capacities ahead of need."""
        while self.pending:
            slot, cap_r, cap_s = self.pending[0]
            nr, ns = int(slot[0]), int(slot[1])
            if nr < 0 or ns < 0:
                if not (wait_all or block_oldest):
                    return
                t0 = time.perf_counter()
                while nr < 0 or ns < 0:
                    if time.perf_counter() - t0 > 120.0:
                        raise RuntimeError("libdibr_b200: list-size report never arrived")
                    nr, ns = int(slot[0]), int(slot[1])
                block_oldest = False
            self.pending.pop(0)
            for i, need, cap in ((0, nr, cap_r), (1, ns, cap_s)):
                if need > 0.8 * caps[i]:
                    caps[i] = int(need * 1.5) + 4096
            if (use_r and nr > cap_r) or (use_s and ns > cap_s):
                self.calibrated = False
                raise RuntimeError(
                    "libdibr_b200: an earlier call of this shape needed %d raster / %d soft list "
                    "entries but was issued with room for %d / %d: its result was incomplete. "
                    "The capacity has been raised; re-run that step (DIBR_B200_WORKSPACE=poll "
                    "verifies every call before it returns)." % (nr, ns, cap_r, cap_s))


_lazy = {}


def check_pending(wait=True):
    """Lazy mode: verify every call issued so far (waits for their reports by default)."""
    for key, st in list(_lazy.items()):
        st.verify(_caps[key], True, True, wait_all=wait)


def _alloc(lib, device, B, F, H, W, cap_r, cap_s, cap_p, knum):
    nbytes = lib.dibr_workspace_bytes_for(B, F, H, W, cap_r, cap_s, cap_p, knum)
    if nbytes < 0:
        _lib.check(int(nbytes), "dibr_workspace_bytes_for")
    return torch.empty(int(nbytes), dtype=torch.uint8, device=device), nbytes


def run(kind, device, B, F, H, W, use_r, use_s, call, knum=0):
    """call(ws_ptr, ws_bytes, cap_r, cap_s, cap_p, needed_ptr) -> rc.  Returns
    (ws, cap_r, cap_s, cap_p) with ws the uint8 workspace tensor the call filled."""
    if SYNC_MODE == "lazy" and not torch.cuda.is_current_stream_capturing():
        lib = _lib.lib()
        key = (device.index, kind, B, F, H, W, knum)
        st = _lazy.get(key)
        if st is None:
            st = _lazy[key] = _Lazy()
            _caps.setdefault(key, initial_caps(B, F, H, W, knum))
        caps = _caps[key]
        st.verify(caps, use_r, use_s)
        if st.calibrated:
            if len(st.pending) >= RING:
                st.verify(caps, use_r, use_s, block_oldest=True)
            use_p = USE_PAIRS and use_s and 0 < knum <= PAIR_MAX_K
            if use_p:
                # what the most recent verified call wanted for its pair buffer
                prev = st.slots[(st.next - 1) % RING]
                want = max(int(prev[2]), int(prev[3]))
                if want > 0.8 * caps[2]:
                    caps[2] = int(want * 1.5) + 1024
            cap_r = caps[0] if use_r else 0
            cap_s = caps[1] if use_s else 0
            cap_p = (caps[2] if PAIR_CAP_OVERRIDE is None else PAIR_CAP_OVERRIDE) if use_p else 0
            cap_p = min(cap_p, 1 << 23)
            ws, nbytes = _alloc(lib, device, B, F, H, W, cap_r, cap_s, cap_p, knum)
            slot = st.slots[st.next]
            st.next = (st.next + 1) % RING
            slot.fill_(-1)
            _lib.check(call(ws.data_ptr(), nbytes, cap_r, cap_s, cap_p, slot.data_ptr()),
                       "libdibr_b200")
            st.pending.append((slot, cap_r, cap_s))
            return ws, cap_r, cap_s, cap_p
        # first call of the shape (or after an overflow): verified before it returns
        out = _run_checked(kind, device, B, F, H, W, use_r, use_s, call, knum, "poll")
        slot = _pinned_slot(key)
        for i in (0, 1):
            caps[i] = max(caps[i], int(int(slot[i]) * 1.5) + 4096)
        st.slots[(st.next - 1) % RING][2:] = slot[2:]
        st.calibrated = True
        return out
    if torch.cuda.is_current_stream_capturing():
        raise RuntimeError("kaolin-windows_b200: this call sizes its workspace from a report the "
                           "device writes to the host and cannot be captured into a CUDA graph; "
                           "use dibr_rasterization(..., cuda_graph=True), which captures the "
                           "library calls itself")
    return _run_checked(kind, device, B, F, H, W, use_r, use_s, call, knum,
                        "sync" if SYNC_MODE == "sync" else "poll")


def _run_checked(kind, device, B, F, H, W, use_r, use_s, call, knum, mode):
    lib = _lib.lib()
    key = (device.index, kind, B, F, H, W, knum)
    caps = _caps.get(key)
    if caps is None:
        caps = _caps[key] = initial_caps(B, F, H, W, knum)
    use_p = USE_PAIRS and use_s and 0 < knum <= PAIR_MAX_K
    if use_p and mode != "sync":
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
        if mode == "sync":
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
