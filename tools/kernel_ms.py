#!/usr/bin/env python
"""Print ms_per_step and the per-kernel times of a bench.py JSON line (stdin)."""
import json
import sys

for line in sys.stdin:
    line = line.strip()
    if not line.startswith("{"):
        continue
    d = json.loads(line)
    k = d["roofline"]["kernel_ms"]
    print("%.4f ms/step | " % d["ms_per_step"] + " ".join("%s=%.1f" % (n, 1e3 * v) for n, v in k.items()))
