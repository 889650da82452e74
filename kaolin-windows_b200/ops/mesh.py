"""kaolin.ops.mesh.index_vertices_by_faces on B200 (kaolin/ops/mesh/mesh.py:25-47): the
vertex -> face-corner gather that feeds the DIB-R path (`face_features`,
`face_vertices_*`), with a scatter-add kernel as its backward instead of torch.gather's
generic index backward.  Same signature, asserts and result layout as the reference.
"""
import ctypes

import torch
from torch.autograd import Function

from .. import _lib

__all__ = ["index_vertices_by_faces"]


def _ptr(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _stream(device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _check_faces(faces, device):
    if faces.dtype != torch.long:
        raise TypeError("faces must be a LongTensor, got %s" % faces.dtype)
    if faces.shape[-1] != 3:
        raise NotImplementedError("kaolin-windows_b200 handles triangle meshes only "
                                  "(faces of shape (num_faces, 3))")
    if faces.device != device:
        raise RuntimeError("faces must be on the same GPU as the vertex tensor")
    return faces.contiguous()


def _check_float(t, name):
    if not t.is_cuda:
        raise RuntimeError("%s must be a CUDA tensor (there is no CPU path)" % name)
    if t.dtype != torch.float32:
        raise TypeError("%s must be float32, got %s" % (name, t.dtype))
    return t.contiguous()


class _IndexVerticesByFaces(Function):
    @staticmethod
    def forward(ctx, vertices_features, faces):
        vf = _check_float(vertices_features, "vertices_features")
        faces = _check_faces(faces, vf.device)
        B, V, C = vf.shape
        F = faces.shape[0]
        out = torch.empty((B, F, 3, C), dtype=torch.float32, device=vf.device)
        with torch.cuda.device(vf.device):
            _lib.check(_lib.lib().dibr_index_vertices_by_faces(
                B, V, F, C, 0, _ptr(vf), _ptr(faces), _ptr(out), _stream(vf.device)),
                "index_vertices_by_faces")
        ctx.save_for_backward(faces)
        ctx.num_vertices = V
        return out

    @staticmethod
    def backward(ctx, grad_out):
        faces, = ctx.saved_tensors
        g = _check_float(grad_out, "grad")
        B, F, _, C = g.shape
        V = ctx.num_vertices
        gv = torch.zeros((B, V, C), dtype=torch.float32, device=g.device)
        with torch.cuda.device(g.device):
            _lib.check(_lib.lib().dibr_scatter_face_grads(
                B, V, F, C, 0, _ptr(g), _ptr(faces), _ptr(gv), C, 0, _stream(g.device)),
                "index_vertices_by_faces backward")
        return gv, None


def index_vertices_by_faces(vertices_features, faces):
    r"""Index vertex features to convert per vertex tensor to per vertex per face tensor.

    Args:
        vertices_features (torch.FloatTensor): of shape (batch_size, num_points, knum).
        faces (torch.LongTensor): face index, of shape (num_faces, 3).

    Returns:
        (torch.FloatTensor): of shape (batch_size, num_faces, 3, knum).
    """
    assert vertices_features.ndim == 3, \
        "vertices_features must have 3 dimensions of shape (batch_size, num_points, knum)"
    assert faces.ndim == 2, "faces must have 2 dimensions of shape (num_faces, num_vertices)"
    return _IndexVerticesByFaces.apply(vertices_features, faces)
