"""TEST INFRASTRUCTURE ONLY -- builds oracle/_ref/kaolin_ref_C.so.

Compiles the reference's own DIB-R source files, unmodified and from where
they lie under /root/reference (nothing is copied into this repo):

    kaolin/csrc/render/mesh/rasterization.cpp
    kaolin/csrc/render/mesh/rasterization_cuda.cu
    kaolin/csrc/render/mesh/dibr_soft_mask.cpp
    kaolin/csrc/render/mesh/dibr_soft_mask_cuda.cu
    kaolin/csrc/render/mesh/deftet.cpp          (next row 8(f).3)
    kaolin/csrc/render/mesh/deftet_cuda.cu

with the reference's own nvcc flags (setup.py:207-211: -O3 -DWITH_CUDA, no
fast-math, default -fmad=true) for sm_100a, plus oracle/ref_binding.cpp, into a
torch extension.  The reference has NO CPU implementation of this path
(rasterization.cpp:100-102 raises without CUDA), so the result only runs on the
GPU box; there it is the authoritative oracle for the -m gpu parity tests and
the "reference kernels on this box" column of bench.py.

Outputs go only to oracle/_ref/ (git-ignored, but shipped by gpurun).
If /root/reference is absent (GPU box) the prebuilt .so is used as is.
"""
import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("KAOLIN_REFERENCE", "/root/reference")
OUT = os.path.join(HERE, "_ref")
SO = os.path.join(OUT, "kaolin_ref_C.so")
# keep the kernels' objects (cuobjdump -sass shows the contraction nvcc chose)
KEEP_OBJS = bool(os.environ.get("KAOLIN_REF_KEEP_OBJS"))
SRC = [
    "kaolin/csrc/render/mesh/rasterization.cpp",
    "kaolin/csrc/render/mesh/dibr_soft_mask.cpp",
    "kaolin/csrc/render/mesh/rasterization_cuda.cu",
    "kaolin/csrc/render/mesh/dibr_soft_mask_cuda.cu",
    # the sibling op on the same math
    "kaolin/csrc/render/mesh/deftet.cpp",
    "kaolin/csrc/render/mesh/deftet_cuda.cu",
]


def _run(cmd):
    print(" ".join(cmd), flush=True)
    subprocess.check_call(cmd)


def build(force=False):
    if not os.path.isdir(REF):
        if os.path.exists(SO):
            return SO
        raise RuntimeError("reference tree %s absent and no prebuilt %s" % (REF, SO))
    srcs = [os.path.join(REF, s) for s in SRC] + [os.path.join(HERE, "ref_binding.cpp")]
    if (not force and os.path.exists(SO)
            and all(os.path.getmtime(SO) > os.path.getmtime(s) for s in srcs)):
        return SO
    import torch
    from torch.utils import cpp_extension as ce
    os.makedirs(OUT, exist_ok=True)
    inc = ["-I" + p for p in ce.include_paths()] + [
        "-I" + sysconfig.get_paths()["include"],
        "-I" + os.path.join(REF, "kaolin", "csrc"),
        "-I/usr/local/cuda/include",
    ]
    defs = ["-DWITH_CUDA", "-DTORCH_EXTENSION_NAME=kaolin_ref_C",
            "-DTORCH_API_INCLUDE_EXTENSION_H", "-D_GLIBCXX_USE_CXX11_ABI=1"]
    objs = []
    procs = []
    for s in srcs:
        o = os.path.join(OUT, os.path.basename(s) + ".o")
        objs.append(o)
        if s.endswith(".cu"):
            cmd = ["nvcc", "-std=c++17", "-O3", "-Xcompiler", "-fPIC",
                   "-DTHRUST_IGNORE_CUB_VERSION_CHECK", "--expt-relaxed-constexpr",
                   "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
                   "-c", s, "-o", o] + inc + defs
        else:
            cmd = ["g++", "-std=c++17", "-O3", "-fPIC", "-c", s, "-o", o] + inc + defs
        print(" ".join(cmd), flush=True)
        procs.append(subprocess.Popen(cmd))
    for p in procs:
        if p.wait() != 0:
            raise RuntimeError("compile failed")
    libdir = ce.library_paths()[0]
    _run(["g++", "-shared", "-o", SO] + objs + [
        "-L" + libdir, "-L/usr/local/cuda/lib64", "-Wl,-rpath," + libdir,
        "-lc10", "-ltorch_cpu", "-ltorch", "-ltorch_python", "-lc10_cuda",
        "-ltorch_cuda", "-lcudart"])
    for o in objs:
        if not (KEEP_OBJS and o.endswith(".cu.o")):
            os.remove(o)
    return SO


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
