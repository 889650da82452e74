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

__all__ = ["prepare_vertices", "texture_mapping"]


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


class _TextureMapping(Function):
    @staticmethod
    def forward(ctx, texture_coordinates, texture_maps, mode):
        coords = _check_float(texture_coordinates, "texture_coordinates")
        tex = _check_float(texture_maps, "texture_maps")
        if tex.device != coords.device:
            raise RuntimeError("texture_mapping: all tensors must be on the same GPU")
        B = coords.shape[0]
        C, Ht, Wt = tex.shape[1:]
        N = coords.numel() // (2 * B) if B > 0 else 0
        out = torch.empty((B, N, C), dtype=torch.float32, device=coords.device)
        with torch.cuda.device(coords.device):
            _lib.check(_lib.lib().dibr_texture_mapping_fwd(
                B, N, C, Ht, Wt, mode, _ptr(coords), _ptr(tex), _ptr(out),
                _stream(coords.device)), "texture_mapping")
        ctx.save_for_backward(coords, tex)
        ctx.mode = mode
        return out

    @staticmethod
    def backward(ctx, grad_out):
        coords, tex = ctx.saved_tensors
        B = coords.shape[0]
        C, Ht, Wt = tex.shape[1:]
        N = coords.numel() // (2 * B) if B > 0 else 0
        g = _check_float(grad_out, "grad")
        g_tex = torch.zeros_like(tex) if ctx.needs_input_grad[1] else None
        g_coords = torch.empty_like(coords) if ctx.needs_input_grad[0] else None
        with torch.cuda.device(coords.device):
            _lib.check(_lib.lib().dibr_texture_mapping_bwd(
                B, N, C, Ht, Wt, ctx.mode, _ptr(coords), _ptr(tex), _ptr(g), _ptr(g_tex),
                _ptr(g_coords), _stream(coords.device)), "texture_mapping backward")
        return g_coords, g_tex, None


def texture_mapping(texture_coordinates, texture_maps, mode='nearest'):
    r"""Interpolates texture_maps by dense or sparse texture_coordinates
    (kaolin/render/mesh/utils.py:23-76): one kernel each way instead of
    clamp / scale / flip / grid_sample / permute / reshape.

    Args:
        texture_coordinates (torch.FloatTensor): (batch_size, h, w, 2) or
            (batch_size, num_points, 2), normalized to [0, 1] (clamped), y bottom to top.
        texture_maps (torch.FloatTensor): (batch_size, num_channels, h', w').
        mode (str): 'nearest' or 'bilinear' as torch.nn.functional.grid_sample.

    Returns:
        (torch.FloatTensor): (batch_size, h, w, num_channels) or
        (batch_size, num_points, num_channels).
    """
    if mode not in ('nearest', 'bilinear'):
        raise ValueError("texture_mapping: mode must be 'nearest' or 'bilinear', got %r" % (mode,))
    if texture_coordinates.shape[-1] != 2 or texture_maps.dim() != 4 \
            or texture_maps.shape[0] != texture_coordinates.shape[0]:
        raise RuntimeError("texture_mapping: expected coordinates (B, ..., 2) and maps (B, C, h', w')")
    out = _TextureMapping.apply(texture_coordinates, texture_maps, 1 if mode == 'bilinear' else 0)
    return out.reshape(texture_coordinates.shape[0], *texture_coordinates.shape[1:-1],
                       texture_maps.shape[1])
