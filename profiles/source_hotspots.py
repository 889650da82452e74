#!/usr/bin/env python
"""Per-source-line share of executed warp instructions and stall samples of one kernel.

    python profiles/source_hotspots.py gpurun_out/x.ncu-rep raster_forward [min_pct] [name substring]

Reads `ncu --page source --print-source cuda,sass --csv` (needs -lineinfo at compile time
and --import-source on at capture time).
"""
import csv
import subprocess
import sys


def main():
    rep, pat = sys.argv[1], sys.argv[2]
    min_pct = float(sys.argv[3]) if len(sys.argv) > 3 else 0.7
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source",
                          "cuda,sass", "--kernel-name", "regex:" + pat],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    must = sys.argv[4] if len(sys.argv) > 4 else ""   # substring the function name must hold
    skip = False
    fname, hdr, lines = "", None, []
    first_fn = None
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            fname = r[1].split("/")[-1]
            continue
        if r[0] == "Function Name":
            skip = must not in r[1]
            if not skip and first_fn is None:
                first_fn = r[1]
            continue
        if skip:
            continue
        if r[0] == "Line No":
            hdr = r
            ie = hdr.index("Instructions Executed")
            sa = hdr.index("# Samples")
            te = hdr.index("Thread Instructions Executed")
            continue
        if hdr is None or r[0] == "":
            continue
        try:
            lines.append((fname, int(r[0]), r[1], float(r[ie]), float(r[sa]), float(r[te])))
        except ValueError:
            pass
    tot_i = sum(l[3] for l in lines) or 1
    tot_s = sum(l[4] for l in lines) or 1
    print("%s\ntotal warp instr %.0f, samples %.0f" % (first_fn, tot_i, tot_s))
    print(" instr%  stall%  thr/instr  file:line  source")
    for f, n, s, i, sm, t in lines:
        if 100 * i / tot_i >= min_pct or 100 * sm / tot_s >= min_pct:
            print("%6.2f  %6.2f  %5.1f  %s:%d  %s" % (100 * i / tot_i, 100 * sm / tot_s,
                                                    t / i if i else 0, f, n, s.strip()[:100]))


if __name__ == "__main__":
    main()
