"""Measured FP32 FMA peak of this GPU (tools/fma_peak.cu).

    python tools/fma_peak.py            # build + run, prints one JSON line

bench.py uses measure() for the denominator of `roofline.secondary`; BASELINE.md 2 records
the number measured on the round's B200.
"""
import ctypes
import json
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "fma_peak.cu")
_SO = os.path.join(_HERE, "_build", "libfma_peak.so")


def build(force=False):
    if not force and os.path.exists(_SO) and os.path.getmtime(_SO) >= os.path.getmtime(_SRC):
        return _SO
    os.makedirs(os.path.dirname(_SO), exist_ok=True)
    subprocess.check_call(["nvcc", "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a",
                           "-lineinfo", "-shared", "-Xcompiler", "-fPIC", "-o", _SO, _SRC])
    return _SO


def measure(iters=1 << 16, reps=5):
    """(TFLOP/s, ms of the best launch) -- needs a GPU."""
    h = ctypes.CDLL(build())
    h.fma_peak_run.restype = ctypes.c_double
    ms = ctypes.c_float(0.0)
    tf = float(h.fma_peak_run(int(iters), int(reps), ctypes.byref(ms)))
    if tf <= 0:
        raise RuntimeError("fma_peak_run failed")
    return tf, float(ms.value)


if __name__ == "__main__":
    tf, ms = measure()
    print(json.dumps({"fp32_fma_tflops": tf, "best_launch_ms": ms,
                      "nominal_tflops_at_1965MHz": 148 * 128 * 2 * 1.965e9 / 1e12}))
