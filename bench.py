#!/usr/bin/env python
"""bench.py -- DIB-R fwd+bwd Mpixels/s (views x H x W) on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    torchrun --nproc-per-node N ... bench.py --gpus N ...

A step = one pass of the hot path over one batch: dibr_rasterization forward, then the
backward driven by fixed random cotangents on both outputs
(loss = (features*G1).sum() + (soft_mask*G2).sum()).

Workload at every N: BASELINE.json configs[1] per GPU (synthetic 69,938-face sphere with
radial noise, 8 views, 512x512, D=3 vertex colours, sigmainv=7000 boxlen=0.02 knum=30).
Weak scaling: rank r renders views [8r, 8r+8) of an 8N-view job over the shared mesh; for
N > 1 the step adds the path's only exchange, a NCCL SUM all-reduce of the shared mesh's
vertex-position/feature gradients (after the local face->vertex scatter-add).

Prints ONE JSON line (rank 0).  `value` = device-resident throughput (CUDA events, max
over ranks, L2 flushed between steps outside the timed spans); `e2e` = the same metric
through the public API from pinned HOST buffers (H2D of every step's inputs and D2H of its
loss inside the timed region; inputs double buffered, the next step's batch prefetched on a
copy stream while this step renders, L2 flush inside the span); `roofline` = dominant kernel vs the measured HBM peak;
`cpu_baseline` = the CPU oracle (a port of the reference kernels; the reference has no CPU
implementation of this path) on a bounded sample, rank 0, N = 1 only.

--impl reference: times that CPU port with all host threads on the same config
(each step = a 64-row band of one view; throughput in pixels/s is band-independent).
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CFG = "cfg2_70k_8x512"
SIGMAINV, BOXLEN, KNUM = 7000, 0.02, 30
KERNELS = ["bin_count", "bin_scan", "bin_fill", "raster_fwd", "softmask_fwd", "raster_bwd",
           "softmask_bwd", "softmask_bwd_dense", "softmask_gather", "softmask_eval",
           "softmask_grad", "softmask_fold", "raster_bwd_finish"]
METRIC = "DIB-R fwd+bwd Mpixels/s (views x H x W)"


_STDOUT_FD = None


def emit(line):
    """Write the result line to the real stdout."""
    sys.stdout.flush()
    if _STDOUT_FD is not None:
        os.dup2(_STDOUT_FD, 1)
    print(json.dumps(line), flush=True)


def _teardown(dist):
    """destroy_process_group with a watchdog: the measurement is printed by now, a stuck
    communicator teardown must not keep the job alive."""
    import threading
    t = threading.Timer(20.0, lambda: os._exit(0))
    t.daemon = True
    t.start()
    try:
        dist.destroy_process_group()
    finally:
        t.cancel()


def algorithmic_bytes(kernel, B, F, P, D):
    """API-level compulsory bytes each kernel touches (SURVEY 8(d) per-unit figures split
    by kernel; DESIGN.md "Algorithmic bytes")."""
    fv, px = B * F, B * P
    return {
        "bin_count": fv * (28 + 24), "bin_fill": fv * (28 + 24), "bin_scan": 0,
        "raster_fwd": fv * (40 + 12 * D) + px * (8 + 4 * D),
        "softmask_fwd": fv * 24 + px * 12, "softmask_gather": fv * 24 + px * 12,
        "softmask_eval": fv * 24, "softmask_fold": px * 4,
        "raster_bwd": fv * (48 + 24 * D) + px * (8 + 4 * D),
        "softmask_bwd": fv * 48 + px * 16, "softmask_grad": fv * 48 + px * 16,
    }.get(kernel, 0)


def total_algorithmic_bytes(B, F, P, D):
    return B * P * (32 + 8 * D) + B * F * (88 + 36 * D)


def hbm_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler(threading.Thread):
    FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag = index, [], set(), False
        self.max_mhz = None

    def run(self):
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        # NVML answers in well under a millisecond (the timed regions last tens of
        # milliseconds; one nvidia-smi call takes longer than that); nvidia-smi is the fallback
        try:
            import pynvml
            pynvml.nvmlInit()
            hd = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(hd, pynvml.NVML_CLOCK_SM))
            get_reasons = getattr(pynvml, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
                pynvml.nvmlDeviceGetCurrentClocksThrottleReasons
            bits = {"sw_power_cap": 0x4, "hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20,
                    "hw_thermal_slowdown": 0x40}
            while not self.stop_flag:
                self.samples.append(float(pynvml.nvmlDeviceGetClockInfo(hd, pynvml.NVML_CLOCK_SM)))
                r = int(get_reasons(hd))
                for n, bit in bits.items():
                    if r & bit:
                        self.reasons.add(n)
                time.sleep(0.002)
            return
        except Exception:
            pass
        while not self.stop_flag:
            try:
                out = subprocess.run(
                    ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.FIELDS,
                     "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5)
                p = [x.strip() for x in out.stdout.strip().split(",")]
                self.samples.append(float(p[0]))
                self.max_mhz = float(p[1])
                for n, v in zip(names, p[2:]):
                    if v.lower().startswith("active"):
                        self.reasons.add(n)
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


def host_threads():
    """Host threads this process may use (torchrun exports OMP_NUM_THREADS=1; the CPU legs
    set their thread count explicitly instead of inheriting it)."""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_reference_step(orc, H, W, z, img, feat, nz, g1, g2):
    _, _, _, ctxs = orc.dibr_rasterization_forward(H, W, z, img, feat, nz, SIGMAINV, BOXLEN, KNUM)
    orc.dibr_rasterization_backward(ctxs, g1, g2)


def reference_arm(args, rank, world):
    """The reference's path on host cores: the CPU port (the reference has no CPU code for
    this path: rasterization.cpp:100-102), all host threads, one FULL view of the 8 per
    step (no row band: the per-view fixed costs of the port are amortised as in a full run)."""
    if rank != 0:
        return
    import torch
    import workloads as wl
    from oracle import dibr_oracle as orc
    cores = host_threads()
    orc.set_threads(cores)
    orc.set_row_band()
    views = wl.CONFIGS[CFG]["views"]
    inp = wl.make_config(CFG, views=views, view_slice=(0, 1))
    H, W = inp["H"], inp["W"]
    z, img, feat, nz = (inp[k] for k in ("face_vertices_z", "face_vertices_image",
                                         "face_features", "face_normals_z"))
    g1 = torch.rand(1, H, W, feat.shape[-1])
    g2 = torch.rand(1, H, W)
    for _ in range(min(args.warmup, 2)):
        cpu_reference_step(orc, H, W, z, img, feat, nz, g1, g2)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_reference_step(orc, H, W, z, img, feat, nz, g1, g2)
    dt = (time.perf_counter() - t0) / args.steps
    assert orc.get_threads() == cores
    val = H * W / dt / 1e6
    sample = "view 0 of %d in full (%d px) per step, %d host threads" % (views, H * W, cores)
    emit({
        "impl": "reference", "metric": METRIC, "value": val, "unit": "Mpixels/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": bench_config(views, int(z.shape[1]), H, W, int(feat.shape[-1])),
        "cpu_baseline": {"value": val, "unit": "Mpixels/s", "cores": cores,
                         "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "Mpixels/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    })


def bench_config(views_per_gpu, F, H, W, D):
    """The workload, identical in both arms (prose lives in the line's `notes`)."""
    return {"workload": CFG, "views_per_gpu": views_per_gpu, "faces": F, "H": H, "W": W, "D": D,
            "sigmainv": SIGMAINV, "boxlen": BOXLEN, "knum": KNUM}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--e2e-chunks", type=int, default=1,
                    help="view chunks the e2e loop overlaps H2D copies and rendering with")
    ap.add_argument("--no-cuda-graph", action="store_true",
                    help="eager launches instead of dibr_rasterization(cuda_graph=True)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-ref-gpu", action="store_true")
    args = ap.parse_args()
    # stdout carries the ONE JSON line: whatever libraries print (NCCL's version banner)
    # goes to stderr -- fd 1 points at stderr until the line is written
    global _STDOUT_FD
    sys.stdout.flush()
    _STDOUT_FD = os.dup(1)
    os.dup2(2, 1)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return reference_arm(args, rank, world)

    import torch
    import torch.distributed as dist
    import workloads as wl
    import kaolin_windows_b200 as kal
    from kaolin_windows_b200 import _lib

    if args.warmup < 3:
        args.warmup = 3
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")   # stdout carries the JSON line only
        import datetime
        dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=120))

    # ---- workload: this rank's 8 views of the 8N-view job over the shared mesh
    views_per_gpu = wl.CONFIGS[CFG]["views"]
    lo, hi = kal.parallel.shard_range(views_per_gpu * world, rank, world)
    inp = wl.make_config(CFG, device=dev, views=views_per_gpu * world, view_slice=(lo, hi))
    H, W = inp["H"], inp["W"]
    z, img, feat, nz = (inp[k] for k in ("face_vertices_z", "face_vertices_image",
                                         "face_features", "face_normals_z"))
    faces = inp["faces"]
    B, F, D = z.shape[0], z.shape[1], feat.shape[-1]
    V = inp["vertices"].shape[1]
    P = H * W
    gen = torch.Generator(device=dev).manual_seed(1234 + rank)
    g1 = torch.rand((B, H, W, D), device=dev, generator=gen)
    g2 = torch.rand((B, H, W), device=dev, generator=gen)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2

    GRAPH = not args.no_cuda_graph
    # (capture_collective=True was measured: 4 % faster at N=2, 45 % slower at N=8)
    exchanger = kal.parallel.MeshGradExchange(inp["faces"], inp["vertices"].shape[1]) \
        if GRAPH else None

    def step(img_in, feat_in, z_in=z, nz_in=nz, exchange=True, with_loss=False, graph=GRAPH):
        img_l = img_in.detach().requires_grad_(True)
        feat_l = feat_in.detach().requires_grad_(True)
        out, mask, idx = kal.render.mesh.dibr_rasterization(H, W, z_in, img_l, feat_l, nz_in,
                                                            SIGMAINV, BOXLEN, KNUM,
                                                            cuda_graph=graph)
        if with_loss:
            loss = (out * g1).sum() + (mask * g2).sum()
            loss.backward()
        else:
            # d loss / d out = g1, d loss / d mask = g2: the same backward without timing
            # torch's multiply/reduce kernels, which are not part of the path
            loss = None
            torch.autograd.backward((out, mask), (g1, g2))
        if world > 1 and exchange:
            # local face -> vertex scatter-add (one kernel per gradient into one flat
            # (V, 2+D) buffer), then ONE all-reduce of the shared mesh's gradients
            if exchanger is not None:
                exchanger(img_l.grad, feat_l.grad)
            else:
                kal.parallel.exchange_mesh_grads(faces, V, img_l.grad, feat_l.grad)
        return loss

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    for _ in range(args.warmup):
        step(img, feat)
    barrier()

    # ---- un-timed cross-check of the exchange (N > 1): the same gradients through our
    # peer-memory all-reduce kernel and through scatter + NCCL all_reduce
    exchange_check = None
    if world > 1:
        img_l = img.detach().requires_grad_(True)
        feat_l = feat.detach().requires_grad_(True)
        out, mask, idx = kal.render.mesh.dibr_rasterization(H, W, z, img_l, feat_l, nz,
                                                            SIGMAINV, BOXLEN, KNUM)
        torch.autograd.backward((out, mask), (g1, g2))
        ex = exchanger if exchanger is not None else kal.parallel.MeshGradExchange(faces, V)
        mine = ex(img_l.grad, feat_l.grad).clone()
        want = kal.parallel.exchange_mesh_grads(faces, V, img_l.grad, feat_l.grad)
        rel = (mine - want).abs().max() / want.abs().max().clamp_min(1e-30)
        bits = mine.contiguous().view(torch.int32).to(torch.int64)
        wts = torch.arange(bits.numel(), device=dev, dtype=torch.int64).view_as(bits) % 65521 + 1
        digest = torch.stack([bits.sum(), (bits * wts).sum()])
        every = [torch.zeros_like(digest) for _ in range(world)]
        dist.all_gather(every, digest)
        relmax = rel.clone()
        dist.all_reduce(relmax, op=dist.ReduceOp.MAX)
        exchange_check = {"mode": ex.mode, "max_rel_diff": float(relmax.item()),
                          "bit_identical_across_ranks": all(torch.equal(e, every[0]) for e in every),
                          "against": "dibr_scatter_face_grads + NCCL all_reduce on the same gradients"}
        del out, mask, idx, img_l, feat_l, mine, want
        barrier()

    # ---- timed: exactly K steps, one CUDA-event pair per step, L2 flushed in between
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
           for _ in range(args.steps)]
    barrier()
    for e0, e1 in evs:
        flush.zero_()
        e0.record()
        step(img, feat)
        e1.record()
    barrier()
    step_ms = [e0.elapsed_time(e1) for e0, e1 in evs]
    total_ms = torch.tensor([sum(step_ms)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)
    ms_per_step = float(total_ms.item()) / args.steps
    value = world * B * P / (ms_per_step * 1e-3) / 1e6

    # ---- e2e: public API from pinned host buffers, H2D of inputs + D2H of the loss per step
    h_in = [t.detach().cpu().pin_memory() for t in (z, img, feat, nz)]
    d_in = [torch.empty_like(t, device=dev) for t in h_in]
    h2d = sum(t.numel() * t.element_size() for t in h_in)

    # The user-level loop a host-fed renderer runs, with the prefetch any data loader does:
    # device inputs are double buffered and step i+1's batch goes over on a copy stream while
    # step i renders (the first step of the region has nothing to hide behind; with
    # --e2e-chunks > 1 its batch goes over in view chunks and rendering starts on the first
    # chunk -- off by default, one odd step only adds captures).  Every step's H2D and
    # the D2H read of its loss are inside the timed region; the loss of step i is read before
    # step i+1 is issued, so at most one batch is in flight ahead of the compute.
    n_chunks = max(1, min(args.e2e_chunks, B))
    bounds = [kal.parallel.shard_range(B, c, n_chunks) for c in range(n_chunks)]
    copy_stream = torch.cuda.Stream(dev)
    slots = [d_in, [torch.empty_like(t) for t in d_in]]
    slot_free = [None, None]                   # event: the slot's last readers are done

    def e2e_copy(slot, chunked):
        """H2D of one batch into `slot` on the copy stream; returns [(lo, hi, event)]."""
        if slot_free[slot] is not None:
            copy_stream.wait_event(slot_free[slot])
        ready = []
        with torch.cuda.stream(copy_stream):
            for lo, hi in (bounds if chunked else [(0, B)]):
                for d, h in zip(slots[slot], h_in):
                    d[lo:hi].copy_(h[lo:hi], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(copy_stream)
                ready.append((lo, hi, ev))
        return ready

    def e2e_compute(slot, ready):
        main = torch.cuda.current_stream(dev)
        d = slots[slot]
        total = None
        gi_out = gf_out = None
        for lo, hi, ev in ready:
            main.wait_event(ev)
            img_l = d[1][lo:hi].requires_grad_(True)
            feat_l = d[2][lo:hi].requires_grad_(True)
            out, mask, idx = kal.render.mesh.dibr_rasterization(
                H, W, d[0][lo:hi], img_l, feat_l, d[3][lo:hi], SIGMAINV, BOXLEN, KNUM,
                cuda_graph=GRAPH)
            # backward of the chunk's loss terms (d loss / d out = G1, d loss / d mask = G2)
            torch.autograd.backward((out, mask), (g1[lo:hi], g2[lo:hi]))
            with torch.no_grad():
                loss = (out * g1[lo:hi]).sum() + (mask * g2[lo:hi]).sum()
                total = loss if total is None else total + loss
            if len(ready) == 1:
                gi_out, gf_out = img_l.grad, feat_l.grad
            else:
                if gi_out is None:
                    gi_out, gf_out = torch.empty_like(d[1]), torch.empty_like(d[2])
                gi_out[lo:hi] = img_l.grad
                gf_out[lo:hi] = feat_l.grad
        if world > 1:
            if exchanger is not None:
                exchanger(gi_out, gf_out)
            else:
                kal.parallel.exchange_mesh_grads(faces, V, gi_out, gf_out)
        done = torch.cuda.Event()
        done.record(main)
        slot_free[slot] = done
        return total

    e2e_trace = [] if os.environ.get("DIBR_BENCH_E2E_TRACE") else None

    def e2e_run(n):
        ready = e2e_copy(0, n_chunks > 1)
        for i in range(n):
            if e2e_trace is not None:
                e2e_trace.append(time.perf_counter())
            nxt = e2e_copy((i + 1) & 1, False) if i + 1 < n else None
            flush.zero_()                      # L2 flush, inside the timed span, not subtracted
            total = e2e_compute(i & 1, ready)
            float(total.item())                # D2H read of this step's loss
            ready = nxt

    e2e_run(6)
    e2e_run(6)
    barrier()
    t0 = time.perf_counter()
    for _ in range(1):
        e2e_run(args.steps)
    torch.cuda.synchronize(dev)
    e2e_t = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
    sampler.stop_flag = True      # clocks sampled across both timed regions (device + e2e)
    e2e_s = max(float(e2e_t.item()), 1e-9) / args.steps
    if e2e_trace is not None and rank == 0:
        ts = e2e_trace[-args.steps:]
        print("e2e step starts (ms):", [round((t - ts[0]) * 1e3, 2) for t in ts], file=sys.stderr)
    e2e_val = world * B * P / e2e_s / 1e6

    if rank != 0:
        if world > 1:
            if exchanger is not None:
                exchanger.release()
            torch.cuda.synchronize(dev)
            _teardown(dist)
        return

    # ---- per-kernel times (CUDA events on the launching stream, a separate pass of K steps)
    lib = _lib.lib()
    lib.dibr_profile_begin()
    for _ in range(args.steps):
        flush.zero_()
        # rank 0 only: no collective in this pass; eager launches so that the library's
        # per-kernel CUDA events bracket each kernel (a graph replay has none)
        step(img, feat, exchange=False, graph=False)
    cap = 64 * args.steps
    ids = (ctypes.c_int * cap)()
    ms = (ctypes.c_float * cap)()
    n = lib.dibr_profile_end(ids, ms, cap)
    per_kernel = {}
    for i in range(n):
        per_kernel.setdefault(KERNELS[ids[i]], []).append(ms[i])
    kernel_ms = {k: sum(v) / len(v) for k, v in per_kernel.items()}
    launches_per_step = n // max(args.steps, 1)
    top = max(kernel_ms, key=kernel_ms.get)
    peak, peak_kind = hbm_peak()
    achieved = algorithmic_bytes(top, B, F, P, D) / (kernel_ms[top] * 1e-3) / 1e9
    roofline = {"bound": "hbm", "kernel": top, "achieved": achieved, "peak": peak,
                "peak_kind": peak_kind, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                "kernel_ms": {k: round(v, 4) for k, v in kernel_ms.items()},
                "kernel_share_of_step": kernel_ms[top] / ms_per_step,
                "path_achieved_gbs": total_algorithmic_bytes(B, F, P, D) / (ms_per_step * 1e-3) / 1e9,
                "path_frac": total_algorithmic_bytes(B, F, P, D) / (ms_per_step * 1e-3) / 1e9 / peak}
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as fh:
            roofline["traffic"] = json.load(fh).get(top)
    except Exception:
        pass

    # ---- secondary roofline: FP32 pipe.  N_pair = (pixel centre, valid face)
    # pairs passing the reference's own box test, x 45 FP32-pipe ops per pair, over the
    # rasterizer's time, against the FMA throughput measured on this GPU (tools/fma_peak.cu)
    try:
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import fma_peak
        fma_tf, _ = fma_peak.measure()
        fma_kind = "measured (tools/fma_peak.cu)"
    except Exception as e:
        fma_tf, fma_kind = 148 * 128 * 2 * 1.965e9 / 1e12, "nominal (%r)" % (e,)
    n_pair = wl.raster_pair_count(H, W, img, nz >= 0, 1000.0)
    if "raster_fwd" in kernel_ms:
        sec_ach = n_pair * 45 / (kernel_ms["raster_fwd"] * 1e-3) / 1e12
        roofline["secondary"] = {
            "bound": "fp32", "kernel": "raster_fwd", "n_pair": n_pair, "ops_per_pair": 45,
            "achieved": sec_ach, "peak": fma_tf, "peak_kind": fma_kind, "unit": "TFLOP/s",
            "frac": sec_ach / fma_tf}

    # ---- the drop-in default: eager launches (no cuda_graph kwarg), same timing rules
    ee = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
          for _ in range(args.steps)]
    for _ in range(3):
        step(img, feat, exchange=False, graph=False)
    torch.cuda.synchronize(dev)
    for e0, e1 in ee:
        flush.zero_()
        e0.record()
        step(img, feat, exchange=False, graph=False)
        e1.record()
    torch.cuda.synchronize(dev)
    eager_ms = sum(e0.elapsed_time(e1) for e0, e1 in ee) / args.steps

    # ---- forward only (SURVEY 8(d) asks for it beside fwd+bwd)
    fe = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
          for _ in range(args.steps)]
    with torch.no_grad():
        for e0, e1 in fe:
            flush.zero_()
            e0.record()
            kal.render.mesh.dibr_rasterization(H, W, z, img, feat, nz, SIGMAINV, BOXLEN, KNUM,
                                               cuda_graph=GRAPH)
            e1.record()
    torch.cuda.synchronize(dev)
    fwd_ms = sum(e0.elapsed_time(e1) for e0, e1 in fe) / args.steps

    line = {
        "metric": METRIC, "value": value, "unit": "Mpixels/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": bench_config(B, F, H, W, D),
        "notes": {"l2": "flushed between steps (256 MiB memset outside the per-step event spans)",
                  "cotangents": "fixed random G1, G2 fed to autograd directly (= backward of "
                                "(features*G1).sum() + (soft_mask*G2).sum()); e2e computes "
                                "and reads back that loss",
                  "launch": "CUDA graphs (dibr_rasterization(cuda_graph=True))" if GRAPH
                            else "eager",
                  "parallelism": "views sharded, %d per GPU" % B,
                  "exchange": ("none (N = 1)" if world == 1 else
                               {"peer": "own one-shot all-reduce kernel over NVLink peer memory",
                                "nccl": "NCCL all-reduce"}.get(
                                    getattr(exchanger, "mode", "nccl"), "NCCL all-reduce"))},
        "clocks": sampler.summary(),
        "e2e": {"value": e2e_val, "unit": "Mpixels/s", "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": 8, "ms_per_step": e2e_s * 1e3,
                "view_chunks": n_chunks,
                "pipeline": "double-buffered inputs: step i+1's H2D on a copy stream during "
                            "step i's render; loss of step i read back before step i+1 is issued"},
        "gpu_launches": launches_per_step * args.steps,
        "fwd_only_ms_per_step": fwd_ms,
        "eager_ms_per_step": eager_ms,
        "exchange_check": exchange_check,
        "roofline": roofline,
    }

    # ---- the reference's own CUDA kernels on this B200 (context, not the baseline arm)
    if not args.no_ref_gpu and world == 1:
        try:
            from oracle import dibr_oracle as orc
            ref = orc.load_ref_ops()
            torch.cuda.synchronize(dev)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            reps = 2
            for i in range(reps + 1):
                if i == 1:
                    e0.record()
                _, _, _, ctxs = orc.dibr_rasterization_forward(H, W, z, img, feat, nz, SIGMAINV,
                                                               BOXLEN, KNUM, ops=ref)
                orc.dibr_rasterization_backward(ctxs, g1, g2, ops=ref)
            e1.record()
            torch.cuda.synchronize(dev)
            rms = e0.elapsed_time(e1) / reps
            line["reference_kernels_on_this_gpu"] = {
                "ms_per_step": rms, "value": B * P / (rms * 1e-3) / 1e6, "unit": "Mpixels/s",
                "what": "reference's unmodified CUDA kernels + torch prep, same inputs"}
        except Exception as e:  # the strengthened checker is optional
            line["reference_kernels_on_this_gpu"] = {"unavailable": repr(e)[:200]}

    # ---- CPU baseline: the oracle port on the host cores, bounded sample (view 0 in full)
    if not args.no_cpu_baseline and world == 1:
        from oracle import dibr_oracle as orc
        cores = host_threads()
        orc.set_threads(cores)
        orc.set_row_band()
        cz, cimg, cfeat, cnz = (t[:1].cpu() for t in (z, img, feat, nz))
        cg1, cg2 = g1[:1].cpu(), g2[:1].cpu()
        t0 = time.perf_counter()
        reps = 0
        while reps < 2 or (time.perf_counter() - t0 < 10.0 and reps < 50):
            cpu_reference_step(orc, H, W, cz, cimg, cfeat, cnz, cg1, cg2)
            reps += 1
        dt = (time.perf_counter() - t0) / reps
        line["cpu_baseline"] = {
            "value": P / dt / 1e6, "unit": "Mpixels/s", "cores": orc.get_threads(),
            "kind": "port",
            "sample": "view 0 of %d in full (%d px) x %d repeats, %d host threads"
                      % (B, P, reps, cores)}
    emit(line)
    if world > 1:
        if exchanger is not None:
            exchanger.release()
        torch.cuda.synchronize(dev)
        _teardown(dist)


if __name__ == "__main__":
    main()
