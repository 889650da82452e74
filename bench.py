#!/usr/bin/env python
"""bench.py -- DIB-R fwd+bwd Mpixels/s (views x H x W) on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    torchrun --nproc-per-node N ... bench.py --gpus N ...

Workload at every N: BASELINE.json configs[1] per GPU (synthetic 69,938-face sphere with
radial noise, 8 views, 512x512, D=3 vertex colours, sigmainv=7000 boxlen=0.02 knum=30).
Weak scaling: rank r renders views [8r, 8r+8) of an 8N-view job over the shared mesh.

A step = one pass of the hot path over one batch, from the MESH: camera transform +
projection, dibr_rasterization forward (rasterize + soft mask), then the backward driven by
fixed random cotangents on both outputs (loss = (features*G1).sum() + (soft_mask*G2).sum(),
SURVEY 8(d)) down to the gradients of the shared mesh's vertices (V,3) and vertex features
(V,D) -- `render.mesh.dibr_rasterization_vertices`, whose kernels read the face corners
through `faces` (no (B,F,3,.) face tensors).  For N > 1 the step adds the path's only
exchange: a SUM all-reduce of those vertex-level gradients.  The same work through the
reference-shaped API on pre-built face tensors (`dibr_rasterization`, graph and eager) is
reported beside it as `dropin`.

Prints ONE JSON line (rank 0).  `value` = device-resident throughput (CUDA events, max
over ranks, L2 flushed between steps outside the timed spans); `e2e` = the same step from
pinned HOST buffers: H2D of the step's inputs (vertices, vertex features, cameras) and D2H
of its results (both gradients and the loss) inside the timed region, L2 flush inside the
span; `roofline` = dominant kernel vs the measured HBM peak, `roofline.secondary` = the
rasterizer's FP32 work vs the measured FMA peak; `cpu_baseline` = the CPU oracle (a port of
the reference kernels; the reference has no CPU implementation of this path) on a bounded
sample, rank 0, N = 1 only.

--impl reference: times that CPU port with all host threads on the same config (each step
= one full view of the 8).
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


L2_BYTES = 126 * 1000 * 1000     # B200 L2 (B200_PROFILING.md)


def step_footprint_bytes(job, B, H, W, D, F):
    """Lower bound of the device bytes ONE mesh-level step touches: the API-level tensors
    (inputs, the three images, both cotangents, both gradients), the per-view vertex images,
    and the workspace parts whose sizes the kernels report (raster records, soft list, pair
    buffer) or that are sized by the problem (exact soft rectangles).  The count matrix, the
    order tables and every re-read are left out."""
    from kaolin_windows_b200.render.mesh import dibr_graph
    V = job["verts"].shape[-2]
    api = sum(job[k].numel() * 4 for k in ("verts", "vfeat", "rot", "trans", "proj"))
    api += job["faces32"].numel() * 4 if "faces32" in job else F * 3 * 4
    images = B * H * W * (D * 4 + 4 + 8)              # features, soft mask, face index
    cot = B * H * W * (D + 1) * 4                     # G1, G2
    grads = (job["verts"].numel() + job["vfeat"].numel()) * 4
    vertex_images = B * V * (2 + 3) * 4
    rect = B * F * 8
    lists = 0
    ents = [e for e in dibr_graph._entries.values() if isinstance(e, dibr_graph._VertexEntry)]
    if ents:
        need = [int(x) for x in ents[-1].slot_np[:4]]     # raster entries, soft entries, pair
        if min(need) >= 0:                                # segments, active regions
            lists = need[0] * 48 + need[1] * 4 + need[2] * 256 * 9 + need[3] * (64 + 4 + 16)
    return int(api + images + cot + grads + vertex_images + rect + lists)


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
    ap.add_argument("--no-cuda-graph", action="store_true",
                    help="eager launches instead of cuda_graph=True")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-ref-gpu", action="store_true")
    ap.add_argument("--no-extras", action="store_true",
                    help="skip the un-timed cfg 3 / cfg 5 sharding runs (`extra`)")
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
    mesh_api = kal.render.mesh.dibr_rasterization_vertices
    GRAPH = not args.no_cuda_graph
    K, Wm = args.steps, args.warmup
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2

    def barrier(coll=True):
        if world > 1 and coll:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def make_mesh_job(cfg, n_views, lo, hi, shared=True):
        """This rank's views [lo, hi) of an n_views job: the mesh (shared by all views, or one
        mesh per view), its cameras, fixed cotangents."""
        inp = wl.make_config(cfg, device=dev, views=n_views, view_slice=(lo, hi))
        rot, trans, proj = inp["camera"]
        verts, vfeat = inp["vertices"].contiguous(), inp["vertex_features"].contiguous()
        if shared:
            verts, vfeat = verts[:1].contiguous(), vfeat[:1].contiguous()
        B = rot.shape[0]
        gen = torch.Generator(device=dev).manual_seed(1234 + rank)
        return dict(inp=inp, H=inp["H"], W=inp["W"], B=B, verts=verts, vfeat=vfeat,
                    faces32=inp["faces"].int().contiguous(), rot=rot.contiguous(),
                    trans=trans.contiguous(), proj=proj.contiguous(),
                    g1=torch.rand((B, inp["H"], inp["W"], vfeat.shape[-1]), device=dev, generator=gen),
                    g2=torch.rand((B, inp["H"], inp["W"]), device=dev, generator=gen))

    def mesh_step(job, exchanger=None, graph=GRAPH, verts=None, vfeat=None):
        """One step from the mesh; returns (loss terms' outputs, vertex grad, feature grad)."""
        v = (job["verts"] if verts is None else verts).detach().requires_grad_(True)
        f = (job["vfeat"] if vfeat is None else vfeat).detach().requires_grad_(True)
        # exchange=: the gradients come back summed over all ranks -- the path's only exchange
        # (N > 1); with graphs it is the last node of the backward graph
        out, mask, idx = mesh_api(job["H"], job["W"], v, job["faces32"], f, job["proj"],
                                  job["rot"], job["trans"], sigmainv=SIGMAINV, boxlen=BOXLEN,
                                  knum=KNUM, cuda_graph=graph, exchange=exchanger)
        # d loss / d out = G1, d loss / d mask = G2: the backward of the loss without timing
        # torch's multiply / reduce kernels, which are not part of the path
        torch.autograd.backward((out, mask), (job["g1"], job["g2"]))
        return out, mask, v.grad, f.grad

    def time_steps(fn, steps, warm, coll=True):
        """ms per step: CUDA events per step, L2 flushed in between, max over ranks
        (coll=False: this rank alone, no collective call)."""
        for _ in range(warm):
            fn()
        barrier(coll)
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
               for _ in range(steps)]
        for e0, e1 in evs:
            flush.zero_()
            e0.record()
            fn()
            e1.record()
        barrier(coll)
        tot = torch.tensor([sum(e0.elapsed_time(e1) for e0, e1 in evs)], device=dev,
                           dtype=torch.float64)
        if world > 1 and coll:
            dist.all_reduce(tot, op=dist.ReduceOp.MAX)
        return float(tot.item()) / steps

    # ---- workload: this rank's 8 views of the 8N-view job over the shared mesh
    views_per_gpu = wl.CONFIGS[CFG]["views"]
    lo, hi = kal.parallel.shard_range(views_per_gpu * world, rank, world)
    job = make_mesh_job(CFG, views_per_gpu * world, lo, hi)
    inp = job["inp"]
    H, W, B = job["H"], job["W"], job["B"]
    z, img, feat, nz = (inp[k] for k in ("face_vertices_z", "face_vertices_image",
                                         "face_features", "face_normals_z"))
    faces = inp["faces"]
    F, D = z.shape[1], feat.shape[-1]
    V = job["verts"].shape[1]
    P = H * W
    g1, g2 = job["g1"], job["g2"]
    exchanger = kal.parallel.VertexGradExchange() if world > 1 else None

    # ---- un-timed cross-check of the exchange (N > 1): the same gradients through our
    # peer-memory all-reduce kernel and through one flat NCCL all_reduce
    exchange_check = None
    if world > 1:
        _, _, gv, gf = mesh_step(job, None, graph=False)
        want = torch.cat([gv.reshape(-1), gf.reshape(-1)]).clone()
        dist.all_reduce(want, op=dist.ReduceOp.SUM)
        for _ in range(3):       # eager first, then captured, then replayed: the timed form
            _, _, xv, xf = mesh_step(job, exchanger)
        mine = torch.cat([xv.reshape(-1), xf.reshape(-1)]).clone()
        rel = (mine - want).abs().max() / want.abs().max().clamp_min(1e-30)
        bits = mine.contiguous().view(torch.int32).to(torch.int64)
        wts = torch.arange(bits.numel(), device=dev, dtype=torch.int64) % 65521 + 1
        digest = torch.stack([bits.sum(), (bits * wts).sum()])
        every = [torch.zeros_like(digest) for _ in range(world)]
        dist.all_gather(every, digest)
        dist.all_reduce(rel, op=dist.ReduceOp.MAX)
        exchange_check = {"mode": exchanger.mode, "max_rel_diff": float(rel.item()),
                          "bit_identical_across_ranks": all(torch.equal(e, every[0]) for e in every),
                          "floats": int(mine.numel()),
                          "folded_into_backward_graph": bool(GRAPH and exchanger.mode == "peer"),
                          "against": "one flat NCCL all_reduce(SUM) of this rank's un-exchanged "
                                     "gradients (eager run of the same step)"}
        del mine, want, gv, gf, xv, xf
        barrier()

    # ---- timed: exactly K steps, one CUDA-event pair per step, L2 flushed in between
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ms_per_step = time_steps(lambda: mesh_step(job, exchanger), K, Wm)
    value = world * B * P / (ms_per_step * 1e-3) / 1e6

    # ---- e2e: the same step from pinned HOST buffers.  Per step, inside the timed region:
    # ONE H2D copy of the step's inputs (vertices, vertex features, cameras, packed in one
    # pinned buffer), the step, and D2H copies of its results (both gradients) into pinned
    # host buffers (`with_loss`: plus the loss, computed by torch).  The user-level loop a
    # host-fed optimiser runs, with the prefetch any data loader does: device inputs are
    # double buffered, step i + 1's H2D and step i's D2H go over a copy stream while the
    # other step computes, and the host reads step i's results while step i + 1 is in flight.
    # Every step's copies and the final read are inside the region.
    names = ("verts", "vfeat", "rot", "trans", "proj")
    sizes = [job[k].numel() for k in names]
    offs, off = [], 0
    for n_k in sizes:                          # every segment 256-byte aligned (the kernels take
        offs.append(off)                       # 16-byte aligned tensors)
        off += (n_k + 63) // 64 * 64
    h_pack = torch.zeros(off, dtype=torch.float32).pin_memory()
    for k, o, n_k in zip(names, offs, sizes):
        h_pack[o:o + n_k] = job[k].detach().reshape(-1).cpu()
    d_packs = [torch.empty_like(h_pack, device=dev) for _ in range(2)]
    e2e_jobs = []
    for d_pack in d_packs:
        d_in = {k: d_pack[o:o + n_k].view(job[k].shape) for k, o, n_k in zip(names, offs, sizes)}
        e2e_jobs.append(dict(job, **d_in))
    nv, nf = job["verts"].numel(), job["vfeat"].numel()
    h_res = [torch.empty(nv + nf + 1, dtype=torch.float32).pin_memory() for _ in range(2)]
    res_ev = [None, None]                      # results of the slot are in host memory
    in_ready = [None, None]                    # H2D into the slot has finished
    in_free = [None, None]                     # the step that read the slot has finished
    keep = [None, None]                        # tensors a pending D2H still reads
    copy_stream = torch.cuda.Stream(dev)
    on_copy_stream = torch.cuda.stream(copy_stream)   # one context object, entered per step
    h2d = h_pack.numel() * 4
    d2h_grads = (nv + nf) * 4
    l2_small = flush[:160 << 20]              # 160 MiB > the 126 MB L2

    def e2e_h2d(slot):
        if in_free[slot] is not None:
            copy_stream.wait_event(in_free[slot])
        with on_copy_stream:
            d_packs[slot].copy_(h_pack, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(copy_stream)
        in_ready[slot] = ev

    def e2e_step(i, with_loss, last, flush_l2):
        main = torch.cuda.current_stream(dev)
        slot = i & 1
        if flush_l2:
            l2_small.zero_()                   # L2 flush, inside the timed span, not subtracted
        main.wait_event(in_ready[slot])
        # (the previous step's D2H reads private copies of its gradients -- result_grads()
        # clones, also when the exchange is folded in -- held in keep[] until that copy has
        # been synchronised: this step does not wait for it)
        j = e2e_jobs[slot]
        out, mask, gv, gf = mesh_step(j, exchanger, verts=j["verts"], vfeat=j["vfeat"])
        if not last:
            e2e_h2d(slot ^ 1)                  # the next step's inputs, behind this step's launch
        loss = None
        if with_loss:
            with torch.no_grad():
                loss = (out * g1).sum() + (mask * g2).sum()
        done = torch.cuda.Event()
        done.record(main)
        in_free[slot] = done
        keep[slot] = (gv, gf, loss)
        copy_stream.wait_event(done)
        h = h_res[slot]
        with on_copy_stream:
            h[:nv].copy_(gv.reshape(-1), non_blocking=True)          # the step's results: both
            h[nv:nv + nf].copy_(gf.reshape(-1), non_blocking=True)   # gradients of the mesh
            if with_loss:
                h[nv + nf:].copy_(loss.reshape(1), non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(copy_stream)
        res_ev[slot] = ev
        # the previous step's results are on the host by now: read them
        prev = res_ev[slot ^ 1]
        if prev is not None:
            prev.synchronize()
            return float(h_res[slot ^ 1][0])
        return None

    def e2e_time(with_loss, flush_l2):
        for n_run in (12, K):
            if n_run == K:
                barrier()
                t0 = time.perf_counter()
            res_ev[0] = res_ev[1] = None
            e2e_h2d(0)
            for i in range(n_run):
                e2e_step(i, with_loss, i == n_run - 1, flush_l2)
            res_ev[(n_run - 1) & 1].synchronize()
            float(h_res[(n_run - 1) & 1][0])   # the last step's results
            torch.cuda.synchronize(dev)
        t = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return max(float(t.item()), 1e-9) / K

    # L2 between e2e steps.  The two input slots are two CUDA-graph entries (graphs are keyed
    # by input addresses), each with its OWN workspace, outputs and gradient buffers, so
    # consecutive steps touch disjoint buffer sets: a buffer is touched again after two steps'
    # worth of other data has gone through the cache.  When one step's footprint alone
    # (counted below from the sizes the kernels report: a lower bound) exceeds the 126 MB L2,
    # that is the timing rules' "inputs larger than L2" and the e2e loop runs no explicit
    # flush; otherwise it keeps the 160 MiB memset inside its timed region.  The flushed
    # variant is always reported beside it (`with_l2_flush`).
    e2e_time(False, True)                      # (settles both slots' graph entries)
    if os.environ.get("DIBR_BENCH_PROFILE") == "1" and rank == 0:   # dev aid: host cost of the loop
        import cProfile
        import pstats
        pr = cProfile.Profile()
        K_keep, K = K, 300
        pr.enable()
        e2e_time(False, False)
        pr.disable()
        K = K_keep
        pstats.Stats(pr, stream=sys.stderr).sort_stats(os.environ.get("DIBR_BENCH_PROFILE_SORT", "tottime")).print_stats(45)
    footprint = step_footprint_bytes(job, B, H, W, D, F)
    e2e_flush = footprint <= L2_BYTES
    e2e_s = e2e_time(False, e2e_flush)
    e2e_loss_s = e2e_time(True, e2e_flush)
    e2e_flushed_s = e2e_s if e2e_flush else e2e_time(False, True)
    sampler.stop_flag = True      # clocks sampled across both timed regions (device + e2e)
    e2e_val = world * B * P / e2e_s / 1e6

    # ---- un-timed by the driver: the sharded configs BASELINE.json names for multi-GPU
    extra = {}
    if not args.no_extras:
        # cfg 3: ONE 200k-face mesh, 64 views at 1024^2, views sharded 64/N per rank (strong
        # scaling) + the vertex-gradient all-reduce
        lo3, hi3 = kal.parallel.shard_range(64, rank, world)
        j3 = make_mesh_job("cfg3_200k_64x1024", 64, lo3, hi3)
        ex3 = kal.parallel.VertexGradExchange() if world > 1 else None   # its own buffers
        ms3 = time_steps(lambda: mesh_step(j3, ex3), 5, 3)
        extra["cfg3_strong"] = {"views_total": 64, "views_per_rank": hi3 - lo3, "H": 1024, "W": 1024,
                                "faces": int(j3["faces32"].shape[0]), "ms_per_step": ms3,
                                "Mpixels_per_s": 64 * 1024 * 1024 / (ms3 * 1e-3) / 1e6,
                                "exchange_floats": int(j3["verts"].numel() + j3["vfeat"].numel())}
        if ex3 is not None:
            ex3.release()
        del j3, ex3
        # cfg 5: 256 distinct 2k-face meshes at 128^2, 256/N per rank, no collective
        lo5, hi5 = kal.parallel.shard_range(256, rank, world)
        j5 = make_mesh_job("cfg5_256x2k_128", 256, lo5, hi5, shared=False)
        ms5 = time_steps(lambda: mesh_step(j5, None), 5, 3)
        extra["cfg5_dp"] = {"meshes_total": 256, "meshes_per_rank": hi5 - lo5, "H": 128, "W": 128,
                            "faces": int(j5["faces32"].shape[0]), "ms_per_step": ms5,
                            "Mpixels_per_s": 256 * 128 * 128 / (ms5 * 1e-3) / 1e6,
                            "collective": "none"}
        del j5
        torch.cuda.empty_cache()

    if rank != 0:
        if world > 1:
            exchanger.release()
            torch.cuda.synchronize(dev)
            _teardown(dist)
        return

    # ---- rank 0 only from here (no collective): per-kernel times, CUDA events on the
    # launching stream around every kernel; eager launches (a graph replay has no events)
    lib = _lib.lib()
    lib.dibr_profile_begin()
    for _ in range(K):
        flush.zero_()
        mesh_step(job, None, graph=False)
    cap = 64 * K
    ids = (ctypes.c_int * cap)()
    ms = (ctypes.c_float * cap)()
    n = lib.dibr_profile_end(ids, ms, cap)
    per_kernel = {}
    for i in range(n):
        per_kernel.setdefault(KERNELS[ids[i]], []).append(ms[i])
    kernel_ms = {k: sum(v) / len(v) for k, v in per_kernel.items()}
    launches_per_step = n // max(K, 1) + 2          # + the two vertex-transform kernels
    top = max(kernel_ms, key=kernel_ms.get)
    peak, peak_kind = hbm_peak()
    achieved = algorithmic_bytes(top, B, F, P, D) / (kernel_ms[top] * 1e-3) / 1e9
    total_bytes = total_algorithmic_bytes(B, F, P, D)
    roofline = {"bound": "hbm", "kernel": top, "achieved": achieved, "peak": peak,
                "peak_kind": peak_kind, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                "kernel_ms": {k: round(v, 4) for k, v in kernel_ms.items()},
                "kernel_share_of_step": kernel_ms[top] / ms_per_step,
                "path_achieved_gbs": total_bytes / (ms_per_step * 1e-3) / 1e9,
                "path_frac": total_bytes / (ms_per_step * 1e-3) / 1e9 / peak,
                "path_bytes": "SURVEY 8(d) face-tensor formula (56 B/px + 196 B/face-view); the "
                              "mesh path's own compulsory traffic is smaller"}
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

    # ---- the same step eagerly launched, and forward only
    eager_ms = time_steps(lambda: mesh_step(job, None, graph=False), K, 3, coll=False)

    def fwd_only():
        with torch.no_grad():
            mesh_api(H, W, job["verts"], job["faces32"], job["vfeat"], job["proj"], job["rot"],
                     job["trans"], sigmainv=SIGMAINV, boxlen=BOXLEN, knum=KNUM, cuda_graph=GRAPH)
    fwd_ms = time_steps(fwd_only, K, 3, coll=False)

    # ---- drop-in API on pre-built face tensors (the reference's call, same pixels)
    def dropin_step(graph):
        img_l = img.detach().requires_grad_(True)
        feat_l = feat.detach().requires_grad_(True)
        out, mask, idx = kal.render.mesh.dibr_rasterization(H, W, z, img_l, feat_l, nz, SIGMAINV,
                                                            BOXLEN, KNUM, cuda_graph=graph)
        torch.autograd.backward((out, mask), (g1, g2))
    dropin = {"api": "dibr_rasterization(face tensors): gradients w.r.t. face_vertices_image "
                     "(B,F,3,2) and face_features (B,F,3,D)",
              "graph_ms_per_step": time_steps(lambda: dropin_step(True), K, 3, coll=False),
              "eager_ms_per_step": time_steps(lambda: dropin_step(False), K, 3, coll=False)}
    dropin["Mpixels_per_s"] = B * P / (dropin["graph_ms_per_step"] * 1e-3) / 1e6

    line = {
        "metric": METRIC, "value": value, "unit": "Mpixels/s", "n_gpus": world,
        "steps": K, "warmup": Wm, "ms_per_step": ms_per_step,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": bench_config(B, F, H, W, D),
        "notes": {"api": "render.mesh.dibr_rasterization_vertices: world-space vertices (V,3), "
                         "faces (F,3), vertex features (V,D), cameras in; features / soft mask / "
                         "face index out; gradients w.r.t. vertices and vertex features back",
                  "l2": "flushed between steps (256 MiB memset outside the per-step event spans)",
                  "cotangents": "fixed random G1, G2 fed to autograd directly (= backward of "
                                "(features*G1).sum() + (soft_mask*G2).sum()); e2e computes "
                                "and reads back that loss",
                  "launch": "CUDA graphs (cuda_graph=True): one launch each way" if GRAPH
                            else "eager",
                  "parallelism": "views sharded, %d per GPU" % B,
                  "exchange": ("none (N = 1)" if world == 1 else
                               {"peer": "own one-shot all-reduce kernel over NVLink peer memory",
                                "nccl": "NCCL all-reduce"}.get(exchanger.mode, "NCCL all-reduce")
                               + " of the vertex-level gradients (V x (3 + D) floats)")},
        "clocks": sampler.summary(),
        "e2e": {"value": e2e_val, "unit": "Mpixels/s", "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h_grads, "ms_per_step": e2e_s * 1e3,
                "what": "per step: one H2D copy of vertices, vertex features and cameras from a "
                        "pinned buffer; the step; D2H of its results (both gradients) into "
                        "pinned buffers. Inputs double buffered: step i+1's H2D and step i's D2H "
                        "on a copy stream; the host reads step i's results while step i+1 runs",
                "l2": ("160 MiB memset inside every timed step" if e2e_flush else
                       "no explicit flush: the two input slots own disjoint workspaces / outputs "
                       "(two graph entries) and one step touches >= %.0f MB > the %d MB L2"
                       % (footprint / 1e6, L2_BYTES // 1000000)),
                "step_footprint_bytes": footprint,
                "with_l2_flush": {"ms_per_step": e2e_flushed_s * 1e3,
                                  "value": world * B * P / e2e_flushed_s / 1e6,
                                  "what": "the same loop with a 160 MiB memset inside every "
                                          "timed step (not subtracted)"},
                "with_loss": {"ms_per_step": e2e_loss_s * 1e3, "d2h_bytes_per_step": d2h_grads + 4,
                              "value": world * B * P / e2e_loss_s / 1e6,
                              "what": "the same + (features*G1).sum() + (soft_mask*G2).sum() "
                                      "computed by torch and read back"}},
        "gpu_launches": launches_per_step * K,
        "fwd_only_ms_per_step": fwd_ms,
        "eager_ms_per_step": eager_ms,
        "dropin": dropin,
        "exchange_check": exchange_check,
        "extra": extra,
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
                "what": "reference's unmodified CUDA kernels + torch prep on face tensors, same pixels"}
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
        exchanger.release()
        torch.cuda.synchronize(dev)
        _teardown(dist)


if __name__ == "__main__":
    main()
