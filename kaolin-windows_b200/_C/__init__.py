"""Reference-shaped raw ops: the same call signatures and return tuples as the pybind11
module ``kaolin._C`` exposes for this path (kaolin/csrc/bindings.cpp:76-81), backed by
libdibr_b200.  Lets the reference's own tests run against the new kernels."""
from . import render
