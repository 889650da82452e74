"""TEST INFRASTRUCTURE ONLY -- CPU oracle of the reference's DIB-R hot path.

Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` may import this module.  The product
package never does (it fails loudly when its CUDA library is missing).

Two layers, both following /root/reference (kaolin 0.14.0):

* raw ops with the signatures of ``kaolin._C.render.mesh.*``
  (kaolin/csrc/bindings.cpp:76-81) -- thin ctypes calls into
  ``oracle/dibr_oracle.cpp`` (the loop-for-loop restatement of the four CUDA
  kernels), including the output allocation/fill of the C++ host wrappers
  (rasterization.cpp:87-93,156-157; dibr_soft_mask.cpp:84-97,172);
* ``rasterize`` / ``dibr_soft_mask`` / ``dibr_rasterization`` forward and
  backward, restating the torch-side preparation of
  kaolin/render/mesh/rasterization.py:299-388 and kaolin/render/mesh/dibr.py:29-73,119-209
  on CPU tensors.

Parity: pinned (see the header of dibr_oracle.cpp and tests/test_oracle_golden.py).
"""
import ctypes
import os
import subprocess

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "dibr_oracle.cpp")
_SO = os.path.join(_HERE, "_build", "libdibr_oracle.so")
_LIB = None

# -ffp-contract=off: the FMA shapes are spelled out in the source (see its header).
CXXFLAGS = ["-O3", "-mavx2", "-mfma", "-ffp-contract=off", "-fopenmp", "-fPIC", "-shared",
            "-std=c++17"]


def build(force=False):
    """Compile oracle/dibr_oracle.cpp -> oracle/_build/libdibr_oracle.so (gcc only)."""
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
        _LIB.oracle_get_threads.restype = ctypes.c_int
    return _LIB


def set_threads(n):
    lib().oracle_set_threads(int(n))


def set_row_band(lo=0, hi=0x7fffffff):
    """Restrict all four kernels to pixel rows [lo, hi) (bench.py's bounded CPU sample)."""
    lib().oracle_set_row_band(int(lo), int(hi))


def get_threads():
    return int(lib().oracle_get_threads())


def _sfx(t):
    if t.dtype == torch.float32:
        return "f32"
    if t.dtype == torch.float64:
        return "f64"
    raise TypeError("oracle not implemented for %s" % t.dtype)


def _p(t):
    assert t.device.type == "cpu" and t.is_contiguous()
    return ctypes.c_void_p(t.data_ptr())


_i, _f = ctypes.c_int, ctypes.c_float


# --------------------------------------------------------------------------- raw ops
def packed_rasterize_forward(height, width, face_vertices_z, face_vertices_image,
                             face_bboxes, face_features, first_idx_face_per_mesh,
                             multiplier, eps, return_pair_count=False):
    """kaolin/csrc/render/mesh/rasterization.cpp:49-104 -> (interp, idx, weights)."""
    z = face_vertices_z.contiguous()
    img = face_vertices_image.contiguous()
    bbox = face_bboxes.contiguous()
    feat = face_features.contiguous()
    first = first_idx_face_per_mesh.contiguous()
    B = first.shape[0] - 1
    D = feat.shape[2]
    idx = torch.full((B, height, width), -1, dtype=torch.long)
    weights = torch.zeros((B, height, width, 3), dtype=z.dtype)
    interp = torch.zeros((B, height, width, D), dtype=z.dtype)
    pairs = ctypes.c_int64(0)
    getattr(lib(), "oracle_packed_rasterize_forward_" + _sfx(z))(
        _p(z), _p(img), _p(bbox), _p(feat), _p(first), _p(idx), _p(weights), _p(interp),
        _i(B), _i(height), _i(width), _i(D), _f(multiplier), _f(eps), ctypes.byref(pairs))
    if return_pair_count:
        return interp, idx, weights, pairs.value
    return interp, idx, weights


def rasterize_backward(grad_interpolated_features, interpolated_features, selected_face_idx,
                       output_weights, face_vertices_image, face_features, eps):
    """rasterization.cpp:106-168 -> (grad_face_vertices_image, grad_face_features)."""
    grad = grad_interpolated_features.contiguous()
    img = face_vertices_image.contiguous()
    feat = face_features.contiguous()
    B, H, W, D = grad.shape
    F = img.shape[1]
    gi = torch.zeros_like(img)
    gf = torch.zeros_like(feat)
    getattr(lib(), "oracle_rasterize_backward_" + _sfx(grad))(
        _p(grad), _p(selected_face_idx.contiguous()), _p(output_weights.contiguous()),
        _p(img), _p(feat), _p(gi), _p(gf), _i(B), _i(H), _i(W), _i(F), _i(D), _f(eps))
    return gi, gf


def dibr_soft_mask_forward(face_vertices_image, face_large_bboxes, selected_face_idx,
                           sigmainv, knum, multiplier):
    """dibr_soft_mask.cpp:48-108 -> (soft_mask, close_face_prob, close_face_idx, dist_type)."""
    img = face_vertices_image.contiguous()
    bbox = face_large_bboxes.contiguous()
    idx = selected_face_idx.contiguous()
    B, F = img.shape[:2]
    H, W = idx.shape[1:]
    cidx = torch.full((B, H, W, knum), -1, dtype=torch.long)
    prob = torch.zeros((B, H, W, knum), dtype=img.dtype)
    typ = torch.zeros((B, H, W, knum), dtype=torch.uint8)
    mask = torch.zeros((B, H, W), dtype=img.dtype)
    getattr(lib(), "oracle_dibr_soft_mask_forward_" + _sfx(img))(
        _p(img), _p(bbox), _p(idx), _p(prob), _p(cidx), _p(typ), _p(mask),
        _i(B), _i(H), _i(W), _i(F), _i(knum), _f(sigmainv), _f(multiplier))
    return mask, prob, cidx, typ


def dibr_soft_mask_backward(grad_soft_mask, soft_mask, selected_face_idx, close_face_prob,
                            close_face_idx, close_face_dist_type, face_vertices_image,
                            sigmainv, multiplier):
    """dibr_soft_mask.cpp:110-183 -> grad_face_vertices_image."""
    img = face_vertices_image.contiguous()
    B, F = img.shape[:2]
    H, W = selected_face_idx.shape[1:]
    K = close_face_idx.shape[-1]
    gi = torch.zeros_like(img)
    getattr(lib(), "oracle_dibr_soft_mask_backward_" + _sfx(img))(
        _p(grad_soft_mask.contiguous()), _p(soft_mask.contiguous()),
        _p(selected_face_idx.contiguous()), _p(close_face_prob.contiguous()),
        _p(close_face_idx.contiguous()), _p(close_face_dist_type.contiguous()), _p(img),
        _p(gi), _i(B), _i(H), _i(W), _i(F), _i(K), _f(sigmainv), _f(multiplier))
    return gi


# ------------------------------------------------------------------ python-level ops
class _CpuOps:
    """The CPU oracle under the raw-op names of kaolin._C.render.mesh."""
    packed_rasterize_forward_cuda = staticmethod(packed_rasterize_forward)
    rasterize_backward_cuda = staticmethod(rasterize_backward)
    dibr_soft_mask_forward_cuda = staticmethod(dibr_soft_mask_forward)
    dibr_soft_mask_backward_cuda = staticmethod(dibr_soft_mask_backward)


def load_ref_ops():
    """The reference's own CUDA kernels (oracle/_ref/kaolin_ref_C.so, built by
    oracle/build_ref.py from /root/reference): same four raw ops, GPU tensors only."""
    import importlib.util
    from . import build_ref
    so = build_ref.build()
    spec = importlib.util.spec_from_file_location("kaolin_ref_C", so)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def rasterize_forward(height, width, face_vertices_z, face_vertices_image, face_features,
                      valid_faces=None, multiplier=None, eps=None, return_pair_count=False,
                      ops=_CpuOps):
    """RasterizeCuda.forward (rasterization.py:289-369) + the list handling and defaults
    of rasterize() (:469-481).  Returns (features, face_idx, ctx) with ctx the saved
    tensors the backward needs."""
    if multiplier is None:
        multiplier = 1000
    if eps is None:
        eps = 1e-8
    if isinstance(face_features, (list, tuple)):
        face_features = torch.cat(face_features, dim=-1)
    batch_size, num_faces = face_vertices_z.shape[:2]
    feat_dim = face_features.shape[-1]
    face_features = face_features.contiguous()
    face_vertices_image = face_vertices_image.contiguous()
    device = face_vertices_z.device
    if valid_faces is None:
        valid_faces_idx = (
            torch.arange(batch_size, device=device).reshape(-1, 1).repeat(1, num_faces).reshape(-1),
            torch.arange(num_faces, device=device).reshape(1, -1).repeat(batch_size, 1).reshape(-1))
        v_img = face_vertices_image.reshape(batch_size * num_faces, 3, 2)
        v_z = face_vertices_z.reshape(batch_size * num_faces, 3)
        v_feat = face_features.reshape(batch_size * num_faces, 3, feat_dim)
        num_faces_per_mesh = torch.full((batch_size,), num_faces, dtype=torch.long,
                                        device=device)
    else:
        valid_faces_idx = torch.where(valid_faces)
        v_img = face_vertices_image[valid_faces_idx[0], valid_faces_idx[1]]
        v_z = face_vertices_z[valid_faces_idx[0], valid_faces_idx[1]]
        v_feat = face_features[valid_faces_idx[0], valid_faces_idx[1]]
        num_faces_per_mesh = torch.sum(valid_faces.reshape(batch_size, -1), dim=1)
    first_idx = torch.zeros(batch_size + 1, dtype=torch.long, device=device)
    torch.cumsum(num_faces_per_mesh, dim=0, out=first_idx[1:])
    v_img = v_img * multiplier                     # :337
    points_min = torch.min(v_img, dim=1)[0]
    points_max = torch.max(v_img, dim=1)[0]
    v_bbox = torch.cat((points_min, points_max), dim=1)
    kw = {"return_pair_count": True} if return_pair_count else {}
    out = ops.packed_rasterize_forward_cuda(
        height, width, v_z.contiguous(), v_img.contiguous(), v_bbox.contiguous(),
        v_feat.contiguous(), first_idx, multiplier, eps, **kw)
    interp, sel, weights = out[:3]
    if valid_faces_idx[1].numel() > 0:
        face_idx = valid_faces_idx[1][
            (sel + first_idx[:-1].reshape(-1, 1, 1)).reshape(-1).clamp(min=0)]
        face_idx = face_idx.reshape(sel.shape).contiguous()
    else:
        face_idx = sel.clone()
    face_idx[sel == -1] = -1
    ctx = dict(interp=interp, face_idx=face_idx, weights=weights,
               face_vertices_image=face_vertices_image, face_features=face_features, eps=eps)
    if return_pair_count:
        ctx["pair_count"] = out[3]
    return interp, face_idx, ctx


def rasterize_backward_from_ctx(ctx, grad_interpolated_features, ops=_CpuOps):
    """RasterizeCuda.backward (rasterization.py:371-388): grads for image coords and
    features only; face_vertices_z gets None."""
    return ops.rasterize_backward_cuda(
        grad_interpolated_features.contiguous(), ctx["interp"], ctx["face_idx"],
        ctx["weights"], ctx["face_vertices_image"], ctx["face_features"], ctx["eps"])


def dibr_soft_mask_forward_py(face_vertices_image, selected_face_idx, sigmainv=7000,
                              boxlen=0.02, knum=30, multiplier=1000., ops=_CpuOps):
    """DibrSoftMaskCuda.forward (dibr.py:29-55).  Returns (soft_mask, ctx)."""
    img = face_vertices_image.contiguous() * multiplier
    sel = selected_face_idx.contiguous()
    points_min = torch.min(img, dim=-2)[0]
    points_max = torch.max(img, dim=-2)[0]
    bbox = torch.cat([points_min - boxlen * multiplier,
                      points_max + boxlen * multiplier], dim=-1)
    mask, prob, cidx, typ = ops.dibr_soft_mask_forward_cuda(img, bbox.contiguous(), sel,
                                                            sigmainv, knum, multiplier)
    ctx = dict(soft_mask=mask, img=img, sel=sel, prob=prob, cidx=cidx, typ=typ,
               sigmainv=sigmainv, multiplier=multiplier)
    return mask, ctx


def dibr_soft_mask_backward_from_ctx(ctx, grad_soft_mask, ops=_CpuOps):
    """DibrSoftMaskCuda.backward (dibr.py:57-73)."""
    return ops.dibr_soft_mask_backward_cuda(
        grad_soft_mask.contiguous(), ctx["soft_mask"], ctx["sel"], ctx["prob"], ctx["cidx"],
        ctx["typ"], ctx["img"], ctx["sigmainv"], ctx["multiplier"])


def dibr_rasterization_forward(height, width, face_vertices_z, face_vertices_image,
                               face_features, face_normals_z, sigmainv=7000, boxlen=0.02,
                               knum=30, multiplier=None, eps=None, ops=_CpuOps):
    """dibr_rasterization (dibr.py:119-209).  Returns (features, soft_mask, face_idx, ctxs)."""
    interp, face_idx, rctx = rasterize_forward(height, width, face_vertices_z,
                                               face_vertices_image, face_features,
                                               face_normals_z >= 0., multiplier, eps, ops=ops)
    _multiplier = 1000. if multiplier is None else multiplier
    mask, sctx = dibr_soft_mask_forward_py(face_vertices_image, face_idx, sigmainv, boxlen,
                                           knum, _multiplier, ops=ops)
    return interp, mask, face_idx, (rctx, sctx)


def dibr_rasterization_backward(ctxs, grad_features, grad_soft_mask, ops=_CpuOps):
    """Sum of the two autograd branches: returns (grad_face_vertices_image,
    grad_face_features)."""
    rctx, sctx = ctxs
    gi, gf = rasterize_backward_from_ctx(rctx, grad_features, ops=ops)
    gi2 = dibr_soft_mask_backward_from_ctx(sctx, grad_soft_mask, ops=ops)
    return gi + gi2, gf


def ctxs_to_double(ctxs):
    """The saved forward state with every floating tensor cast to float64 and every
    discrete choice (face indices, nearest-feature types) kept: running the backward on it
    gives a low-noise reference for the SAME discrete decisions the fp32 forward took."""
    def lift(ctx):
        return {k: (v.double() if torch.is_tensor(v) and v.is_floating_point() else v)
                for k, v in ctx.items()}
    return tuple(lift(c) for c in ctxs)


def mask_iou(lhs_mask, rhs_mask):
    """kaolin/metrics/render.py:18-40 (used by the reference's golden-gradient tests)."""
    batch_size = lhs_mask.shape[0]
    sil_mul = lhs_mask * rhs_mask
    sil_add = lhs_mask + rhs_mask
    iou_up = torch.sum(sil_mul.reshape(batch_size, -1), dim=1)
    iou_down = torch.sum((sil_add - sil_mul).reshape(batch_size, -1), dim=1)
    iou_neg = iou_up / (iou_down + 1e-10)
    return 1.0 - torch.mean(iou_neg)
