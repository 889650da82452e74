// TEST INFRASTRUCTURE ONLY -- not part of the product path.
//
// pybind11 module exposing the UNMODIFIED reference DIB-R host wrappers
// (compiled from the sources where they lie under /root/reference, see
// oracle/build_ref.py) under the same names kaolin/csrc/bindings.cpp:76-81
// registers them with, so the -m gpu parity tests can call the reference's own
// CUDA kernels on the B200 box.  The four hot-path ops and the two deftet ops
// (bindings.cpp:82-83) are bound.
#include <torch/extension.h>
#include "render/mesh/rasterization.h"
#include "render/mesh/dibr_soft_mask.h"
#include "render/mesh/deftet.h"

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.def("packed_rasterize_forward_cuda", &kaolin::packed_rasterize_forward_cuda);
  m.def("rasterize_backward_cuda", &kaolin::rasterize_backward_cuda);
  m.def("dibr_soft_mask_forward_cuda", &kaolin::dibr_soft_mask_forward_cuda);
  m.def("dibr_soft_mask_backward_cuda", &kaolin::dibr_soft_mask_backward_cuda);
  m.def("deftet_sparse_render_forward_cuda", &kaolin::deftet_sparse_render_forward_cuda);
  m.def("deftet_sparse_render_backward_cuda", &kaolin::deftet_sparse_render_backward_cuda);
}
