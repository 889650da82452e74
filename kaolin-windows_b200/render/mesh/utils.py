"""kaolin.render.mesh.prepare_vertices on B200 (kaolin/render/mesh/utils.py:128-175): the
step right before the DIB-R path.  The camera transform and the projection work on
(batch, vertices)-sized tensors and stay in torch (so gradients reach the camera
parameters exactly as in the reference); everything (batch, faces)-sized -- the two
index_vertices_by_faces gathers and the unit face normals, forward and backward -- is one
CUDA kernel each way (csrc/dibr_mesh.cu).
"""
import ctypes

import torch
from torch.autograd import Function

from ... import _lib
from ...ops.mesh import _check_faces, _check_float, _ptr, _stream

__all__ = ["prepare_vertices"]


def _rotate_translate_points(points, camera_rot, camera_trans):
    # kaolin/render/camera/legacy.py:22-37
    translated_points = points - camera_trans.view(-1, 1, 3)
    return torch.matmul(translated_points, camera_rot.permute(0, 2, 1))


def _perspective_camera(points, camera_proj):
    # kaolin/render/camera/legacy.py:120-139
    projected_points = points * camera_proj.view(-1, 1, 3)
    return projected_points[:, :, :2] / projected_points[:, :, 2:3]


class _FaceSetup(Function):
    @staticmethod
    def forward(ctx, vertices_camera, vertices_image, faces):
        vc = _check_float(vertices_camera, "vertices_camera")
        vi = _check_float(vertices_image, "vertices_image")
        faces = _check_faces(faces, vc.device)
        B, V, _ = vc.shape
        F = faces.shape[0]
        dev = vc.device
        fvc = torch.empty((B, F, 3, 3), dtype=torch.float32, device=dev)
        fvi = torch.empty((B, F, 3, 2), dtype=torch.float32, device=dev)
        fn = torch.empty((B, F, 3), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().dibr_face_setup_fwd(
                B, V, F, _ptr(vc), _ptr(vi), _ptr(faces), _ptr(fvc), _ptr(fvi), _ptr(fn),
                _stream(dev)), "prepare_vertices")
        ctx.save_for_backward(vc, faces)
        ctx.set_materialize_grads(False)
        return fvc, fvi, fn

    @staticmethod
    def backward(ctx, g_fvc, g_fvi, g_fn):
        vc, faces = ctx.saved_tensors
        B, V, _ = vc.shape
        F = faces.shape[0]
        dev = vc.device
        g_fvc, g_fvi, g_fn = (None if g is None else _check_float(g, "grad")
                              for g in (g_fvc, g_fvi, g_fn))
        g_vc = g_vi = None
        if g_fvc is not None or g_fn is not None:
            g_vc = torch.zeros((B, V, 3), dtype=torch.float32, device=dev)
        if g_fvi is not None:
            g_vi = torch.zeros((B, V, 2), dtype=torch.float32, device=dev)
        if g_vc is not None or g_vi is not None:
            with torch.cuda.device(dev):
                _lib.check(_lib.lib().dibr_face_setup_bwd(
                    B, V, F, _ptr(vc), _ptr(faces), _ptr(g_fvc), _ptr(g_fvi), _ptr(g_fn),
                    _ptr(g_vc), _ptr(g_vi), _stream(dev)), "prepare_vertices backward")
        return g_vc, g_vi, None


def prepare_vertices(vertices, faces, camera_proj, camera_rot=None, camera_trans=None,
                     camera_transform=None):
    r"""Wrapper function to move and project vertices to cameras then index them with faces.

    Args:
        vertices (torch.Tensor): the meshes vertices, of shape (batch_size, num_vertices, 3).
        faces (torch.LongTensor): the meshes faces, of shape (num_faces, 3).
        camera_proj (torch.Tensor): the camera projection vector, of shape (3, 1).
        camera_rot (torch.Tensor, optional): rotation matrices, of shape (batch_size, 3, 3).
        camera_trans (torch.Tensor, optional): translation vectors, of shape (batch_size, 3).
        camera_transform (torch.Tensor, optional): transformation matrices, of shape
            (batch_size, 4, 3). Replace `camera_trans` and `camera_rot`.

    Returns:
        (torch.Tensor, torch.Tensor, torch.Tensor):
            face_vertices_camera (batch_size, num_faces, 3, 3),
            face_vertices_image (batch_size, num_faces, 3, 2),
            unit face normals (batch_size, num_faces, 3).
    """
    if camera_transform is None:
        assert camera_trans is not None and camera_rot is not None, \
            "camera_transform or camera_trans and camera_rot must be defined"
        vertices_camera = _rotate_translate_points(vertices, camera_rot, camera_trans)
    else:
        assert camera_trans is None and camera_rot is None, \
            "camera_trans and camera_rot must be None when camera_transform is defined"
        padded_vertices = torch.nn.functional.pad(vertices, (0, 1), mode='constant', value=1.)
        vertices_camera = padded_vertices @ camera_transform
    vertices_image = _perspective_camera(vertices_camera, camera_proj)
    return _FaceSetup.apply(vertices_camera, vertices_image, faces)
