#!/usr/bin/env python
"""Per-kernel counts of the SASS mnemonics DESIGN.md's kernel descriptions rest on
(static counts from `cuobjdump -sass` of the built library; no GPU needed).

    python profiles/sass_mnemonics.py [kaolin-windows_b200/libdibr_b200.so] > profiles/r2_sass_mnemonics.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "kaolin-windows_b200", "libdibr_b200.so")
# mnemonic prefix -> what it evidences
WATCH = [
    ("MATCH", "__match_any_sync: lanes grouped by face (raster backward)"),
    ("REDUX", "__reduce_*_sync: warp min / max / add in one instruction"),
    ("VOTE", "ballots: bit transposes of the gather, ranks of the ordered fill"),
    ("SHFL", "shuffle reductions / scans"),
    ("ATOMS", "shared-memory atomics: histograms, lane masks, 64-bit z-buffer CAS"),
    ("REDG", "global reductions without return, all forms (gradient accumulation)"),
    ("REDG.E.ADD.F32x2", "of which two-float vector reductions (red.global.add.v2.f32)"),
    ("ATOMG", "global atomics WITH return (cursor reservations only)"),
    ("MUFU.RCP", "reciprocal seeds of the IEEE divisions"),
    ("DADD", "FP64 add (eps of the normalisation, soft-mask arithmetic)"),
    ("DMUL", "FP64 multiply"),
    ("DFMA", "FP64 fused multiply-add"),
    ("LDG.E.128", "16-byte global loads (face records, int4 scan rows)"),
    ("STG.E.128", "16-byte global stores (face records)"),
    ("UBLKCP", "TMA bulk copies (none at HEAD: DESIGN.md 3c)"),
    ("BAR.SYNC", "CTA barriers"),
]
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
fn, counts, total = None, collections.defaultdict(collections.Counter), collections.Counter()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        fn = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        fn = re.sub(r"\(.*", "", fn).replace("void ", "").replace("dibr::", "").replace("(anonymous namespace)::", "")
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Za-z0-9_.]+)", line)
    if fn and m:
        total[fn] += 1
        op = m.group(1)
        for key, _ in WATCH:
            if op.startswith(key):
                counts[fn][key] += 1
print("# static SASS mnemonic counts per kernel of libdibr_b200.so (sm_100a), cuobjdump -sass")
for key, why in WATCH:
    print("#   %-10s %s" % (key, why))
keys = [k for k, _ in WATCH]
print("%-52s %6s " % ("kernel", "instr") + " ".join("%9s" % k for k in keys))
for fn in sorted(total, key=lambda f: -total[f]):
    if total[fn] < 40:
        continue
    print("%-52s %6d " % (fn[:52], total[fn]) + " ".join("%9d" % counts[fn][k] for k in keys))
