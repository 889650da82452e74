"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel.
usage: python profiles/summarize_launches.py gpurun_out/launches.csv > profiles/<name>.txt"""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[1:]:
    t = float(r[vi].replace(",", ""))
    t = t / 1e3 if r[ui] == "ns" else (t * 1e3 if r[ui] == "ms" else t)
    name = r[ki].replace("void ", "")
    agg[name[:70]][0] += 1
    agg[name[:70]][1] += t
tot = sum(v[1] for v in agg.values())
print("# per-kernel device time (cold-cache, serialised under ncu: compare SHARES)")
print("%-72s %6s %12s %7s" % ("kernel", "n", "total_us", "share"))
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("%-72s %6d %12.1f %6.1f%%" % (k, v[0], v[1], 100 * v[1] / tot))
print("%-72s %6s %12.1f" % ("TOTAL", "", tot))
# the library's own kernels only (bench.py also runs torch fills, the reference kernels and
# its FMA-peak micro-kernel, all outside the timed step)
own = {k: v for k, v in agg.items() if k.startswith("dibr::")}
tot_own = sum(v[1] for v in own.values()) or 1.0
print()
print("# libdibr_b200 kernels only: share of the step's own kernel time")
print("%-72s %6s %12s %10s %7s" % ("kernel", "n", "total_us", "us/launch", "share"))
for k, v in sorted(own.items(), key=lambda kv: -kv[1][1]):
    print("%-72s %6d %12.1f %10.1f %6.1f%%" % (k, v[0], v[1], v[1] / v[0], 100 * v[1] / tot_own))
