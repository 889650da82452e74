#!/usr/bin/env python
"""Summarise an `ncu --set full` report: the metrics that matter per path kernel, and
`profiles/traffic.json` (DRAM bytes per launch per kernel, read by bench.py's roofline).

    python profiles/summarize_full.py gpurun_out/x.ncu-rep "header text" > profiles/<name>.txt
"""
import collections
import csv
import json
import os
import subprocess
import sys

METRICS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "smsp__thread_inst_executed_per_inst_executed.ratio",
    "sm__cycles_active.avg", "sm__cycles_active.max", "sm__cycles_elapsed.max",
    "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    # L2 atomic / reduction traffic of the backward kernels (north_star: "L2 atomic throughput")
    "lts__t_sectors_op_atom.sum", "lts__t_sectors_op_red.sum",
    "lts__t_sector_op_red_hit_rate.pct", "lts__t_sector_op_atom_hit_rate.pct",
    "smsp__inst_executed_op_global_red.sum", "smsp__inst_executed_op_shared_atom.sum",
]
# added to `ncu --set full` on the command line (not all are part of the full set)
EXTRA_NCU_METRICS = ",".join(METRICS[-7:])
# bench.py's kernel names by substring of the ncu kernel name (first match wins)
BENCH_NAMES = [
    ("bin_walk_kernel<0>", "bin_count"), ("bin_walk_kernel<false>", "bin_count"),
    ("bin_scan", "bin_scan"), ("bin_walk_kernel", "bin_fill"),
    ("raster_forward", "raster_fwd"),
    ("raster_backward", "raster_bwd"), ("softmask_gather", "softmask_gather"),
    ("softmask_pairs_eval", "softmask_eval"), ("softmask_pairs_grad", "softmask_grad"),
]
SCALE = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def main():
    rep = sys.argv[1]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True,
                         text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    ki = hdr.index("Kernel Name")
    print("# %s; per launch" % (sys.argv[2] if len(sys.argv) > 2 else rep))
    seen = collections.Counter()
    traffic = {}
    for r in rows[2:]:
        name = r[ki]
        seen[name] += 1
        if seen[name] > 1:          # one launch per kernel is enough here
            continue
        print("----\nKernel Name = %s" % name)
        dram = 0.0
        for m in METRICS:
            if m not in hdr:
                continue
            j = hdr.index(m)
            print("%s = %s %s" % (m, r[j], units[j]))
            if m.startswith("dram__bytes"):
                dram += float(r[j].replace(",", "")) * SCALE.get(units[j], 1)
        if "gpu__time_duration.sum" in hdr:
            # derived: L2 reduction/atomic sectors per microsecond of the launch
            try:
                j = hdr.index("gpu__time_duration.sum")
                t_us = float(r[j].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(units[j], 1.0)
                n = sum(float(r[hdr.index(m)].replace(",", "")) for m in
                        ("lts__t_sectors_op_atom.sum", "lts__t_sectors_op_red.sum") if m in hdr)
                if n:
                    print("derived: L2 atom+red sectors per us = %.1f (32 B each: %.1f GB/s)"
                          % (n / t_us, n * 32 / t_us / 1e3))
            except (ValueError, ZeroDivisionError):
                pass
        for sub, bname in BENCH_NAMES:
            if sub in name:
                if bname:
                    traffic[bname] = traffic.get(bname, 0) + dram
                break
    here = os.path.dirname(os.path.abspath(__file__))
    with open(os.path.join(here, "traffic.json"), "w") as fh:
        json.dump({k: int(v) for k, v in traffic.items()}, fh, indent=1)


if __name__ == "__main__":
    main()
