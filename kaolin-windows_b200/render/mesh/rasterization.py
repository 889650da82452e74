"""kaolin.render.mesh.rasterize(..., backend='cuda') on B200.

Same public signature, defaults and return layout as the reference
(kaolin/render/mesh/rasterization.py:390-506); the work of RasterizeCuda
(:289-388: valid-face packing, x multiplier, bounding boxes, the _C kernel, index
remapping) happens inside libdibr_b200's kernels.
"""
import ctypes

import torch
from torch.autograd import Function

from ... import _lib, _workspace

__all__ = ["rasterize"]


def _ptr(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _prep(t, name, dtype=torch.float32):
    """contiguous, 16-byte aligned CUDA tensor of the dtype the kernels take."""
    if not t.is_cuda:
        raise RuntimeError("%s must be a CUDA tensor (there is no CPU path)" % name)
    if t.dtype != dtype:
        if t.dtype == torch.float64 and dtype == torch.float32:
            raise TypeError("%s is float64 while face_vertices_z is float32: all floating "
                            "inputs must share one dtype" % name)
        raise TypeError("%s must be %s, got %s" % (name, dtype, t.dtype))
    t = t.contiguous()
    if t.data_ptr() % 16:
        t = t.clone()
    return t


def _stream(device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


class RasterizeB200(Function):
    """Drop-in for RasterizeCuda (rasterization.py:260-388)."""

    @staticmethod
    def forward(ctx, height, width, face_vertices_z, face_vertices_image, face_features,
                valid_faces, multiplier, eps):
        z = _prep(face_vertices_z, "face_vertices_z")
        img = _prep(face_vertices_image, "face_vertices_image")
        feat = _prep(face_features, "face_features")
        B, F = z.shape[0], z.shape[1]
        D = feat.shape[-1]
        if img.shape != (B, F, 3, 2) or z.shape != (B, F, 3) or feat.shape != (B, F, 3, D):
            raise RuntimeError("rasterize: expected face_vertices_z (B,F,3), face_vertices_image "
                               "(B,F,3,2), face_features (B,F,3,D); got %s %s %s"
                               % (tuple(z.shape), tuple(img.shape), tuple(feat.shape)))
        dev = z.device
        if img.device != dev or feat.device != dev:
            raise RuntimeError("rasterize: all tensors must be on the same GPU")
        valid = None
        if valid_faces is not None:
            if valid_faces.shape != (B, F):
                raise RuntimeError("valid_faces must be of shape (batch_size, num_faces)")
            valid = valid_faces.to(device=dev, dtype=torch.bool).contiguous().view(torch.uint8)
        height, width = int(height), int(width)
        idx = torch.empty((B, height, width), dtype=torch.long, device=dev)
        out = torch.empty((B, height, width, D), dtype=torch.float32, device=dev)
        lib = _lib.lib()
        with torch.cuda.device(dev):
            st = _stream(dev)

            def call(ws, ws_bytes, cap_r, cap_s, cap_p, needed):
                return lib.dibr_rasterize_fwd(
                    B, F, height, width, D, _ptr(z), _ptr(img), _ptr(feat), _ptr(valid), None,
                    0, None, None, float(multiplier), float(eps), _ptr(idx), _ptr(out), None,
                    ws, ws_bytes, cap_r, needed, st)

            _workspace.run("raster", dev, B, F, height, width, True, False, call)
        ctx.save_for_backward(idx, img, feat)
        ctx.mark_non_differentiable(idx)
        ctx.set_materialize_grads(False)
        ctx.eps = float(eps)
        ctx.multiplier = float(multiplier)
        return out, idx

    @staticmethod
    def backward(ctx, grad_interpolated_features, grad_face_idx):
        if grad_interpolated_features is None:
            return (None,) * 8
        idx, img, feat = ctx.saved_tensors
        B, F = img.shape[:2]
        H, W = idx.shape[1:]
        D = feat.shape[-1]
        grad = _prep(grad_interpolated_features, "grad_interpolated_features")
        grad_img = torch.zeros_like(img)
        grad_feat = torch.zeros_like(feat)
        with torch.cuda.device(img.device):
            rc = _lib.lib().dibr_rasterize_bwd(
                B, F, H, W, D, _ptr(grad), _ptr(idx), None, _ptr(img), _ptr(feat),
                ctx.multiplier, ctx.eps, _ptr(grad_img), _ptr(grad_feat), _stream(img.device))
        _lib.check(rc, "dibr_rasterize_bwd")
        # face_vertices_z gets no gradient, as in the reference (rasterization.py:387-388)
        return None, None, None, grad_img, grad_feat, None, None, None


def rasterize(height, width, face_vertices_z, face_vertices_image, face_features,
              valid_faces=None, multiplier=None, eps=None, backend='cuda'):
    r"""Fully differentiable DIB-R rasterization (see the reference docstring,
    kaolin/render/mesh/rasterization.py:399-468, for the argument semantics).

    Args:
        height, width (int): size of the rendered images.
        face_vertices_z (torch.FloatTensor): (batch_size, num_faces, 3).
        face_vertices_image (torch.FloatTensor): (batch_size, num_faces, 3, 2), in [-1, 1].
        face_features (torch.FloatTensor or list): (batch_size, num_faces, 3, feature_dim[i]).
        valid_faces (torch.BoolTensor): (batch_size, num_faces) faces to rasterize. Default: all.
        multiplier (int): coordinates are enlarged by it internally. Default: 1000.
        eps (float): epsilon of the barycentric normalisation. Default: 1e-8.
        backend (str): only 'cuda' (the hand-written sm_100a kernels) exists here; the
            reference's 'nvdiffrast' / 'nvdiffrast_fwd' need OpenGL and are not provided.

    Returns:
        (torch.FloatTensor or tuple, torch.LongTensor): interpolated features
        (batch_size, height, width, feature_dim) -- a tuple if a list was given -- and the
        rendered face index (batch_size, height, width), -1 for background.
    """
    if multiplier is None:
        multiplier = 1000
    if eps is None:
        eps = 1e-8
    _face_features = torch.cat(face_features, dim=-1) \
        if isinstance(face_features, (list, tuple)) else face_features
    if backend == 'cuda':
        fn = RasterizeB200
        if face_vertices_z.dtype == torch.float64:     # the reference dispatches both dtypes
            from ._f64 import RasterizeF64 as fn
        image_features, face_idx = fn.apply(
            height, width, face_vertices_z, face_vertices_image, _face_features, valid_faces,
            multiplier, eps)
    elif backend in ('nvdiffrast', 'nvdiffrast_fwd'):
        raise ValueError('backend "%s" needs nvdiffrast/OpenGL, which kaolin-windows_b200 does '
                         'not provide; use backend="cuda"' % backend)
    else:
        raise ValueError(f'"{backend}" is not a valid backend, '
                         'valid choices are ["cuda", "nvdiffrast", "nvdiffrast_fwd"]')
    if isinstance(face_features, (list, tuple)):
        _image_features = []
        cur_idx = 0
        for face_feature in face_features:
            _image_features.append(image_features[..., cur_idx:cur_idx + face_feature.shape[-1]])
            cur_idx += face_feature.shape[-1]
        image_features = tuple(_image_features)
    return image_features, face_idx
