"""kaolin.render.mesh.deftet_sparse_render on B200.

Same public signature, defaults and return layout as the reference
(kaolin/render/mesh/deftet.py:325-417).  Everything DeftetSparseRenderer.forward does in
torch around its raw op (:273-323: bounding boxes, the _C kernel, argsort by depth, three
gathers, w2, the padded feature gather and the weighted sum) is one call into libdibr_b200
(csrc/dibr_deftet.cu); the backward is the other.  float32 and float64.
"""
import torch
from torch.autograd import Function

from ... import _lib
from .rasterization import _ptr, _stream

__all__ = ["deftet_sparse_render"]

_DTYPES = {torch.float32: 0, torch.float64: 1}     # DIBR_F32 / DIBR_F64


def _prep_like(t, name, dtype, dev):
    if not t.is_cuda:
        raise RuntimeError("%s must be a CUDA tensor (there is no CPU path)" % name)
    if t.device != dev:
        raise RuntimeError("deftet_sparse_render: all tensors must be on the same GPU")
    if t.dtype != dtype:
        raise TypeError("%s must be %s like face_vertices_z, got %s" % (name, dtype, t.dtype))
    t = t.detach().contiguous()
    if t.data_ptr() % 16:
        t = t.clone()
    return t


def workspace(dev, B, F, P):
    n = _lib.lib().dibr_deftet_workspace_bytes(B, F, P)
    if n < 0:
        _lib.check(int(n), "dibr_deftet_workspace_bytes")
    return torch.empty(int(n), dtype=torch.uint8, device=dev), int(n)


class DeftetSparseRendererB200(Function):
    """Drop-in for DeftetSparseRenderer (deftet.py:269-340)."""

    @staticmethod
    def forward(ctx, pixel_coords, render_ranges, face_vertices_z, face_vertices_image,
                face_features, knum, eps):
        dtype = face_vertices_z.dtype
        if dtype not in _DTYPES:
            raise RuntimeError("deftet_sparse_render not implemented for '%s'" % dtype)
        if not face_vertices_z.is_cuda:
            raise RuntimeError("face_vertices_z must be a CUDA tensor (there is no CPU path)")
        dev = face_vertices_z.device
        z = _prep_like(face_vertices_z, "face_vertices_z", dtype, dev)
        img = _prep_like(face_vertices_image, "face_vertices_image", dtype, dev)
        feat = _prep_like(face_features, "face_features", dtype, dev)
        pix = _prep_like(pixel_coords, "pixel_coords", dtype, dev)
        rng = _prep_like(render_ranges, "render_ranges", dtype, dev)
        B, F = z.shape[0], z.shape[1]
        D = feat.shape[-1]
        P = pix.shape[1]
        K = int(knum)
        if (z.shape != (B, F, 3) or img.shape != (B, F, 3, 2) or feat.shape != (B, F, 3, D)
                or pix.shape != (B, P, 2) or rng.shape != (B, P, 2)):
            raise RuntimeError(
                "deftet_sparse_render: expected pixel_coords, render_ranges (B,P,2), "
                "face_vertices_z (B,F,3), face_vertices_image (B,F,3,2), face_features "
                "(B,F,3,D); got %s %s %s %s %s" % (tuple(pix.shape), tuple(rng.shape),
                                                   tuple(z.shape), tuple(img.shape),
                                                   tuple(feat.shape)))
        face_idx = torch.empty((B, P, K), dtype=torch.long, device=dev)
        weights = torch.empty((B, P, K, 3), dtype=dtype, device=dev)
        out = torch.empty((B, P, K, D), dtype=dtype, device=dev)
        with torch.cuda.device(dev):
            ws, ws_bytes = workspace(dev, B, F, P)
            rc = _lib.lib().dibr_deftet_fwd(
                _DTYPES[dtype], B, F, P, K, D, _ptr(z), _ptr(img), _ptr(feat), _ptr(pix),
                _ptr(rng), float(eps), _ptr(face_idx), _ptr(out), _ptr(weights), _ptr(ws),
                ws_bytes, _stream(dev))
        _lib.check(rc, "dibr_deftet_fwd")
        ctx.save_for_backward(face_idx, weights, img, feat)
        ctx.mark_non_differentiable(face_idx)
        ctx.set_materialize_grads(False)
        ctx.eps = float(eps)
        return out, face_idx

    @staticmethod
    def backward(ctx, grad_interpolated_features, grad_face_idx):
        if grad_interpolated_features is None:
            return (None,) * 7
        face_idx, weights, img, feat = ctx.saved_tensors
        B, P, K = face_idx.shape
        F, D = img.shape[1], feat.shape[-1]
        grad = _prep_like(grad_interpolated_features, "grad_interpolated_features", img.dtype,
                          img.device)
        grad_img = torch.empty_like(img)
        grad_feat = torch.empty_like(feat)
        with torch.cuda.device(img.device):
            rc = _lib.lib().dibr_deftet_bwd(
                _DTYPES[img.dtype], B, F, P, K, D, _ptr(grad), _ptr(face_idx), _ptr(weights),
                _ptr(img), _ptr(feat), ctx.eps, _ptr(grad_img), _ptr(grad_feat),
                _stream(img.device))
        _lib.check(rc, "dibr_deftet_bwd")
        # no gradient for pixel_coords, render_ranges, face_vertices_z (deftet.py:340)
        return None, None, None, grad_img, grad_feat, None, None


def deftet_sparse_render(pixel_coords, render_ranges, face_vertices_z,
                         face_vertices_image, face_features, knum=300, eps=1e-8):
    r"""Fully differentiable volumetric renderer of *Gao et al.*, "Learning Deformable
    Tetrahedral Meshes for 3D Reconstruction" (see the reference docstring,
    kaolin/render/mesh/deftet.py:327-397): renders, for every given image coordinate, ALL
    the intersected faces (up to ``knum``) in depth order.

    Args:
        pixel_coords (torch.Tensor): image coordinates to render, (batch_size, num_pixels, 2).
        render_ranges (torch.Tensor): depth ranges [min, max) that get rendered,
            (batch_size, num_pixels, 2).
        face_vertices_z (torch.Tensor): (batch_size, num_faces, 3), higher = closer.
        face_vertices_image (torch.Tensor): (batch_size, num_faces, 3, 2).
        face_features (torch.Tensor or list of torch.Tensor):
            (batch_size, num_faces, 3, feature_dim[i]).
        knum (int): maximum number of faces that influence one pixel. Default: 300.
        eps (float): epsilon of the barycentric normalisation. Default: 1e-8.

    Returns:
        (torch.Tensor or tuple of torch.Tensor, torch.LongTensor): rendered features
        (batch_size, num_pixels, knum, feature_dim) -- a tuple if a list was given -- and the
        rendered face index (batch_size, num_pixels, knum), -1 is void.
    """
    _face_features = torch.cat(
        face_features, dim=-1
    ) if isinstance(face_features, (list, tuple)) else face_features

    image_features, face_idx = DeftetSparseRendererB200.apply(
        pixel_coords, render_ranges, face_vertices_z,
        face_vertices_image, _face_features, knum, eps)
    if isinstance(face_features, (list, tuple)):
        _image_features = []
        cur_idx = 0
        for face_feature in face_features:
            _image_features.append(image_features[..., cur_idx:cur_idx + face_feature.shape[-1]])
            cur_idx += face_feature.shape[-1]
        image_features = tuple(_image_features)
    return image_features, face_idx
