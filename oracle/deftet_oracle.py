"""TEST INFRASTRUCTURE ONLY -- CPU oracle of the reference's deftet_sparse_render.

Only ``tests/`` may import this module; the product package never does.

* raw ops with the signatures of ``kaolin._C.render.mesh.deftet_sparse_render_*``
  (kaolin/csrc/bindings.cpp:82-83): ctypes calls into ``oracle/deftet_oracle.cpp``;
* ``deftet_sparse_render_forward`` / ``_backward``: the torch side of
  DeftetSparseRenderer (kaolin/render/mesh/deftet.py:273-340) restated on CPU tensors.
  The reference's ``torch.argsort(descending=True)`` leaves the order of equal depths
  unspecified; here it is stable (insertion = ascending face order), which is also what
  the product does.

Parity: pinned (header of deftet_oracle.cpp, tests/test_deftet_golden.py).
"""
import ctypes
import os
import subprocess

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "deftet_oracle.cpp")
_SO = os.path.join(_HERE, "_build", "libdeftet_oracle.so")
_LIB = None
CXXFLAGS = ["-O2", "-mfma", "-ffp-contract=off", "-fopenmp", "-fPIC", "-shared", "-std=c++17"]


def build(force=False):
    if (not force and os.path.exists(_SO)
            and os.path.getmtime(_SO) >= os.path.getmtime(_SRC)):
        return _SO
    os.makedirs(os.path.dirname(_SO), exist_ok=True)
    tmp = _SO + ".tmp%d" % os.getpid()
    subprocess.check_call(["g++"] + CXXFLAGS + ["-o", tmp, _SRC])
    os.replace(tmp, _SO)
    return _SO


def lib():
    global _LIB
    if _LIB is None:
        try:
            _LIB = ctypes.CDLL(build())
        except OSError:
            _LIB = ctypes.CDLL(build(force=True))
    return _LIB


def _sfx(t):
    return {torch.float32: "f32", torch.float64: "f64"}[t.dtype]


def _p(t):
    return ctypes.c_void_p(t.data_ptr())


def deftet_sparse_render_forward_cuda(face_vertices_z, face_vertices_image, face_bboxes,
                                      pixel_coords, pixel_depth_ranges, knum, eps):
    """deftet.cpp:47-107 + deftet_cuda.cu:31-193 on CPU tensors."""
    z, img, bbox, pix, rng = (t.contiguous() for t in (
        face_vertices_z, face_vertices_image, face_bboxes, pixel_coords, pixel_depth_ranges))
    B, F = z.shape[:2]
    P = pix.shape[1]
    ids = torch.empty((B, P, knum), dtype=torch.long)
    depth = torch.empty((B, P, knum), dtype=z.dtype)
    w0 = torch.empty((B, P, knum), dtype=z.dtype)
    w1 = torch.empty((B, P, knum), dtype=z.dtype)
    getattr(lib(), "deftet_forward_" + _sfx(z))(
        _p(z), _p(img), _p(bbox), _p(pix), _p(rng), _p(ids), _p(depth), _p(w0), _p(w1),
        B, F, P, int(knum), ctypes.c_float(eps))
    return ids, depth, w0, w1


def deftet_sparse_render_backward_cuda(grad, face_idx, weights, face_vertices_image,
                                       face_features, eps):
    """deftet.cpp:109-165 + deftet_cuda.cu:239-401 on CPU tensors."""
    grad, face_idx, weights, img, feat = (t.contiguous() for t in (
        grad, face_idx, weights, face_vertices_image, face_features))
    B, P, K, D = grad.shape
    F = img.shape[1]
    grad_img = torch.empty_like(img)
    grad_feat = torch.empty_like(feat)
    getattr(lib(), "deftet_backward_" + _sfx(img))(
        _p(grad), _p(face_idx), _p(weights), _p(img), _p(feat), _p(grad_img), _p(grad_feat),
        B, F, P, K, D, ctypes.c_float(eps))
    return grad_img, grad_feat


def deftet_sparse_render_forward(pixel_coords, render_ranges, face_vertices_z,
                                 face_vertices_image, face_features, knum=300, eps=1e-8):
    """DeftetSparseRenderer.forward (deftet.py:273-323).  Returns (features, face_idx,
    weights) -- weights is what the reference saves for its backward."""
    B, F = face_vertices_z.shape[:2]
    D = face_features.shape[-1]
    P = pixel_coords.shape[1]
    face_min = torch.min(face_vertices_image, dim=2)[0]
    face_max = torch.max(face_vertices_image, dim=2)[0]
    face_bboxes = torch.cat((face_min, face_max), dim=2)
    face_idx, pixel_depth, w0, w1 = deftet_sparse_render_forward_cuda(
        face_vertices_z, face_vertices_image, face_bboxes, pixel_coords, render_ranges,
        knum, eps)
    sorted_idx = torch.argsort(pixel_depth, descending=True, dim=-1, stable=True)
    sorted_face_idx = torch.gather(face_idx, -1, sorted_idx).contiguous()
    sorted_w0 = torch.gather(w0, -1, sorted_idx)
    sorted_w1 = torch.gather(w1, -1, sorted_idx)
    sorted_w2 = (sorted_face_idx != -1).float() - (sorted_w0 + sorted_w1)
    _idx = (sorted_face_idx + 1).reshape(B, -1, 1, 1).expand(B, P * knum, 3, D)
    selected = torch.gather(
        torch.nn.functional.pad(face_features, (0, 0, 0, 0, 1, 0), value=0.), 1, _idx
    ).reshape(B, P, knum, 3, D)
    weights = torch.stack([sorted_w0, sorted_w1, sorted_w2], dim=-1).contiguous()
    feats = torch.sum(weights.unsqueeze(-1) * selected, dim=-2).contiguous()
    return feats, sorted_face_idx, weights


def deftet_sparse_render_backward(grad, face_idx, weights, face_vertices_image, face_features,
                                  eps=1e-8):
    """DeftetSparseRenderer.backward (deftet.py:325-340)."""
    return deftet_sparse_render_backward_cuda(grad, face_idx, weights, face_vertices_image,
                                              face_features, eps)
