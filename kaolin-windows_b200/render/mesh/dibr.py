"""kaolin.render.mesh.dibr_soft_mask / dibr_rasterization on B200.

Same public signatures and defaults as kaolin/render/mesh/dibr.py:75-209.  The fast path
keeps no per-pixel face lists between forward and backward (the reference saves
13*knum bytes per pixel, dibr.py:51-54): the backward re-walks the binned face lists left
in the workspace by the forward.
"""
import numpy as np
import torch
from torch.autograd import Function

from ... import _lib, _workspace
from .rasterization import _prep, _ptr, _stream, rasterize

__all__ = ["dibr_soft_mask", "dibr_rasterization", "dibr_rasterization_mesh",
           "dibr_rasterization_vertices"]


def _margin(boxlen, multiplier):
    # dibr.py:37-38: the Python product boxlen * multiplier (double) becomes a float32
    # scalar when subtracted from / added to the float32 box corners
    return float(np.float32(boxlen * multiplier))


class DibrSoftMaskB200(Function):
    """Drop-in for DibrSoftMaskCuda (dibr.py:27-73)."""

    @staticmethod
    def forward(ctx, face_vertices_image, selected_face_idx, sigmainv, boxlen, knum, multiplier):
        img = _prep(face_vertices_image, "face_vertices_image")
        sel = _prep(selected_face_idx, "selected_face_idx", torch.long)
        B, F = img.shape[:2]
        if img.shape != (B, F, 3, 2) or sel.dim() != 3 or sel.shape[0] != B:
            raise RuntimeError("dibr_soft_mask: expected face_vertices_image (B,F,3,2) and "
                               "selected_face_idx (B,H,W)")
        if sel.device != img.device:
            raise RuntimeError("dibr_soft_mask: all tensors must be on the same GPU")
        H, W = sel.shape[1:]
        dev = img.device
        mask = torch.empty((B, H, W), dtype=torch.float32, device=dev)
        margin = _margin(boxlen, multiplier)
        lib = _lib.lib()
        with torch.cuda.device(dev):
            st = _stream(dev)

            def call(ws, ws_bytes, cap_r, cap_s, cap_p, needed):
                return lib.dibr_softmask_fwd(
                    B, F, H, W, _ptr(img), 0, None, _ptr(sel), float(sigmainv), margin, int(knum),
                    float(multiplier), _ptr(mask), None, None, None, ws, ws_bytes, cap_r, cap_s,
                    cap_p, needed, st)

            ws, cap_r, cap_s, cap_p = _workspace.run("soft", dev, B, F, H, W, False, True, call,
                                                     int(knum))
        ctx.save_for_backward(img, sel, mask, ws)
        ctx.args = (float(sigmainv), margin, int(knum), float(multiplier), cap_r, cap_s, cap_p)
        return mask

    @staticmethod
    def backward(ctx, grad_soft_mask):
        img, sel, mask, ws = ctx.saved_tensors
        sigmainv, margin, knum, multiplier, cap_r, cap_s, cap_p = ctx.args
        B, F = img.shape[:2]
        H, W = sel.shape[1:]
        grad = _prep(grad_soft_mask, "grad_soft_mask")
        grad_img = torch.zeros_like(img)
        with torch.cuda.device(img.device):
            rc = _lib.lib().dibr_softmask_bwd(
                B, F, H, W, _ptr(img), 0, None, _ptr(sel), sigmainv, margin, knum, multiplier,
                _ptr(grad), _ptr(mask), _ptr(grad_img), _ptr(ws), ws.numel(), cap_r, cap_s,
                cap_p, _stream(img.device))
        _lib.check(rc, "dibr_softmask_bwd")
        return grad_img, None, None, None, None, None


def dibr_soft_mask(face_vertices_image, selected_face_idx,
                   sigmainv=7000, boxlen=0.02, knum=30, multiplier=1000.):
    r"""Soft silhouette mask of DIB-R (see kaolin/render/mesh/dibr.py:77-115).

    Args:
        face_vertices_image (torch.Tensor): (batch_size, num_faces, 3, 2), in [-1, 1].
        selected_face_idx (torch.LongTensor): (batch_size, height, width), from rasterize.
        sigmainv (float): smoothness term, the higher the sharper. Default: 7000.
        boxlen (float): margin over the face bounding boxes. Default: 0.02.
        knum (int): maximum number of faces that can influence one pixel. Default: 30.
        multiplier (float): internal coordinate multiplier. Default: 1000.

    Returns:
        (torch.FloatTensor): the soft mask, of shape (batch_size, height, width).
    """
    fn = DibrSoftMaskB200
    if face_vertices_image.dtype == torch.float64:     # the reference dispatches both dtypes
        from ._f64 import DibrSoftMaskF64 as fn
    return fn.apply(face_vertices_image, selected_face_idx, sigmainv, boxlen, knum, multiplier)


class DibrRasterizationB200(Function):
    """rasterize(valid = normal_z >= 0) + dibr_soft_mask fused: one binning pass feeds both
    kernels, and one backward accumulates both gradient paths into the same buffer."""

    @staticmethod
    def forward(ctx, height, width, face_vertices_z, face_vertices_image, face_features,
                face_normals_z, sigmainv, boxlen, knum, multiplier, eps):
        z = _prep(face_vertices_z, "face_vertices_z")
        img = _prep(face_vertices_image, "face_vertices_image")
        feat = _prep(face_features, "face_features")
        nz = _prep(face_normals_z, "face_normals_z")
        B, F = z.shape[:2]
        D = feat.shape[-1]
        if (img.shape != (B, F, 3, 2) or z.shape != (B, F, 3) or feat.shape != (B, F, 3, D)
                or nz.shape != (B, F)):
            raise RuntimeError("dibr_rasterization: inconsistent input shapes")
        dev = z.device
        if any(t.device != dev for t in (img, feat, nz)):
            raise RuntimeError("dibr_rasterization: all tensors must be on the same GPU")
        H, W = int(height), int(width)
        idx = torch.empty((B, H, W), dtype=torch.long, device=dev)
        out = torch.empty((B, H, W, D), dtype=torch.float32, device=dev)
        mask = torch.empty((B, H, W), dtype=torch.float32, device=dev)
        margin = _margin(boxlen, multiplier)
        lib = _lib.lib()
        with torch.cuda.device(dev):
            st = _stream(dev)

            def call(ws, ws_bytes, cap_r, cap_s, cap_p, needed):
                return lib.dibr_render_fwd(
                    B, F, H, W, D, _ptr(z), _ptr(img), _ptr(feat), _ptr(nz), float(sigmainv),
                    margin, int(knum), float(multiplier), float(eps), _ptr(idx), _ptr(out),
                    _ptr(mask), ws, ws_bytes, cap_r, cap_s, cap_p, needed, st)

            ws, cap_r, cap_s, cap_p = _workspace.run("render", dev, B, F, H, W, True, True, call,
                                                     int(knum))
        ctx.save_for_backward(idx, img, feat, mask, ws)
        ctx.mark_non_differentiable(idx)
        ctx.set_materialize_grads(False)
        ctx.args = (float(sigmainv), margin, int(knum), float(multiplier), float(eps), cap_r, cap_s,
                    cap_p)
        return out, mask, idx

    @staticmethod
    def backward(ctx, grad_features, grad_soft_mask, grad_face_idx):
        idx, img, feat, mask, ws = ctx.saved_tensors
        sigmainv, margin, knum, multiplier, eps, cap_r, cap_s, cap_p = ctx.args
        B, F = img.shape[:2]
        H, W = idx.shape[1:]
        D = feat.shape[-1]
        if grad_features is None and grad_soft_mask is None:
            return (None,) * 11
        # zero-filled by the call -- with ONE memset, the two being exactly adjacent
        pool = torch.empty(img.numel() + feat.numel(), dtype=torch.float32, device=img.device)
        grad_img = pool[:img.numel()].view(img.shape)
        grad_feat = pool[img.numel():].view(feat.shape)
        g = None if grad_features is None else _prep(grad_features, "grad_interpolated_features")
        gm = None if grad_soft_mask is None else _prep(grad_soft_mask, "grad_soft_mask")
        with torch.cuda.device(img.device):
            # both backward halves (and the zero-fill of the gradients) in one call
            rc = _lib.lib().dibr_render_bwd(
                B, F, H, W, D, _ptr(g), _ptr(gm), _ptr(idx), _ptr(img), _ptr(feat), _ptr(mask),
                sigmainv, margin, knum, multiplier, eps, _ptr(grad_img), _ptr(grad_feat),
                _ptr(ws), ws.numel(), cap_r, cap_s, cap_p, _stream(img.device))
        _lib.check(rc, "dibr_render_bwd")
        return (None, None, None, grad_img, grad_feat, None, None, None, None, None, None)


def dibr_rasterization(height, width, face_vertices_z, face_vertices_image,
                       face_features, face_normals_z, sigmainv=7000,
                       boxlen=0.02, knum=30, multiplier=None, eps=None,
                       rast_backend='cuda', cuda_graph=False):
    r"""Fully differentiable DIB-R renderer (see kaolin/render/mesh/dibr.py:123-188).

    Renders per-vertex per-face features to feature images, a soft foreground mask and a
    face index map.  Only faces with ``face_normals_z >= 0`` are rasterized; the soft mask
    uses all faces (dibr.py:190-208).

    ``cuda_graph=True`` (not in the reference; see dibr_graph.py) replays the forward and
    the backward from CUDA graphs once the same problem (shapes, parameters and input
    addresses) is seen a second time; the results then live in static buffers that the
    next such call overwrites.

    Returns:
        (torch.Tensor or tuple, torch.Tensor, torch.LongTensor): features
        (batch_size, height, width, num_features), soft mask (batch_size, height, width),
        face index (batch_size, height, width), -1 is None.
    """
    if rast_backend != 'cuda':
        # same dispatch (and the same errors) as rasterize()
        interpolated_features, face_idx = rasterize(
            height, width, face_vertices_z, face_vertices_image, face_features,
            face_normals_z >= 0., multiplier, eps, rast_backend)
    if face_vertices_z.dtype == torch.float64 and rast_backend == 'cuda':
        # double: the reference's own composition of its two ops (dibr.py:190-208)
        interpolated_features, face_idx = rasterize(
            height, width, face_vertices_z, face_vertices_image, face_features,
            face_normals_z >= 0., multiplier, eps, rast_backend)
        soft_mask = dibr_soft_mask(face_vertices_image, face_idx, sigmainv, boxlen, knum,
                                   1000. if multiplier is None else multiplier)
        return interpolated_features, soft_mask, face_idx
    _multiplier = 1000. if multiplier is None else multiplier
    _eps = 1e-8 if eps is None else eps
    _face_features = torch.cat(face_features, dim=-1) \
        if isinstance(face_features, (list, tuple)) else face_features
    res = None
    if cuda_graph and rast_backend == 'cuda':
        from .dibr_graph import graphed_dibr_rasterization
        res = graphed_dibr_rasterization(
            height, width, face_vertices_z, face_vertices_image, _face_features, face_normals_z,
            (float(sigmainv), _margin(boxlen, _multiplier), int(knum), float(_multiplier),
             float(_eps)))
    if res is None:
        res = DibrRasterizationB200.apply(
            height, width, face_vertices_z, face_vertices_image, _face_features, face_normals_z,
            sigmainv, boxlen, knum, _multiplier, _eps)
    image_features, soft_mask, face_idx = res
    if isinstance(face_features, (list, tuple)):
        _image_features = []
        cur_idx = 0
        for face_feature in face_features:
            _image_features.append(image_features[..., cur_idx:cur_idx + face_feature.shape[-1]])
            cur_idx += face_feature.shape[-1]
        image_features = tuple(_image_features)
    return image_features, soft_mask, face_idx


# --------------------------------------------------------------------------- mesh level
def _prep_faces(faces, num_vertices):
    """(F, 3) int32 CUDA copy of the face list (cached on the tensor's storage address)."""
    if faces.dim() != 2 or faces.shape[1] != 3:
        raise RuntimeError("faces must be of shape (num_faces, 3)")
    if not faces.is_cuda:
        raise RuntimeError("faces must be a CUDA tensor (there is no CPU path)")
    if faces.dtype == torch.int32 and faces.is_contiguous():
        return faces
    key = (faces.data_ptr(), faces._version, tuple(faces.shape), faces.dtype)
    hit = _faces32_cache.get(key)
    if hit is None:
        if len(_faces32_cache) > 16:
            _faces32_cache.clear()
        hit = _faces32_cache[key] = (faces, faces.to(torch.int32).contiguous())
    return hit[1]


_faces32_cache = {}


class DibrRasterizationMeshB200(Function):
    """dibr_rasterization straight from the mesh: the kernels fetch the face corners through
    `faces`, so prepare_vertices' (B, F, 3, .) face tensors (render/mesh/utils.py:128-175)
    never exist, and the backward adds straight into vertex-level gradients."""

    @staticmethod
    def forward(ctx, height, width, vertices_camera, vertices_image, faces32, vertex_features,
                valid_faces, sigmainv, boxlen, knum, multiplier, eps):
        vcam = _prep(vertices_camera, "vertices_camera")
        vimg = _prep(vertices_image, "vertices_image")
        vfeat = _prep(vertex_features, "vertex_features")
        B, V = vcam.shape[:2]
        F = faces32.shape[0]
        D = vfeat.shape[-1]
        if (vcam.shape != (B, V, 3) or vimg.shape != (B, V, 2) or vfeat.dim() != 3
                or vfeat.shape[1] != V or vfeat.shape[0] not in (1, B)):
            raise RuntimeError("dibr_rasterization_mesh: expected vertices_camera (B,V,3), "
                               "vertices_image (B,V,2), vertex_features (B or 1,V,D)")
        shared = int(vfeat.shape[0] == 1)
        dev = vcam.device
        if any(t.device != dev for t in (vimg, vfeat, faces32)):
            raise RuntimeError("dibr_rasterization_mesh: all tensors must be on the same GPU")
        valid = None
        if valid_faces is not None:
            if valid_faces.shape != (B, F):
                raise RuntimeError("valid_faces must be of shape (batch_size, num_faces)")
            valid = valid_faces.to(device=dev, dtype=torch.bool).contiguous().view(torch.uint8)
        H, W = int(height), int(width)
        idx = torch.empty((B, H, W), dtype=torch.long, device=dev)
        out = torch.empty((B, H, W, D), dtype=torch.float32, device=dev)
        mask = torch.empty((B, H, W), dtype=torch.float32, device=dev)
        margin = _margin(boxlen, multiplier)
        lib = _lib.lib()
        with torch.cuda.device(dev):
            st = _stream(dev)

            def call(ws, ws_bytes, cap_r, cap_s, cap_p, needed):
                return lib.dibr_render_mesh_fwd(
                    B, V, F, H, W, D, _ptr(vcam), _ptr(vimg), _ptr(faces32), _ptr(vfeat), shared,
                    _ptr(valid), float(sigmainv), margin, int(knum), float(multiplier),
                    float(eps), _ptr(idx), _ptr(out), _ptr(mask), ws, ws_bytes, cap_r, cap_s,
                    cap_p, needed, st)

            ws, cap_r, cap_s, cap_p = _workspace.run("mesh", dev, B, F, H, W, True, True, call,
                                                     int(knum))
        ctx.save_for_backward(idx, vimg, vfeat, faces32, mask, ws)
        ctx.mark_non_differentiable(idx)
        ctx.set_materialize_grads(False)
        ctx.args = (float(sigmainv), margin, int(knum), float(multiplier), float(eps), cap_r, cap_s,
                    cap_p, shared)
        return out, mask, idx

    @staticmethod
    def backward(ctx, grad_features, grad_soft_mask, grad_face_idx):
        idx, vimg, vfeat, faces32, mask, ws = ctx.saved_tensors
        sigmainv, margin, knum, multiplier, eps, cap_r, cap_s, cap_p, shared = ctx.args
        B, V = vimg.shape[:2]
        F = faces32.shape[0]
        H, W = idx.shape[1:]
        D = vfeat.shape[-1]
        if grad_features is None and grad_soft_mask is None:
            return (None,) * 12
        lib = _lib.lib()
        # zero-filled by the call -- with ONE memset, the two being exactly adjacent
        pool = torch.empty(vimg.numel() + vfeat.numel(), dtype=torch.float32, device=vimg.device)
        grad_vimg = pool[:vimg.numel()].view(vimg.shape)
        grad_vfeat = pool[vimg.numel():].view(vfeat.shape)
        g = None if grad_features is None else _prep(grad_features, "grad_interpolated_features")
        gm = None if grad_soft_mask is None else _prep(grad_soft_mask, "grad_soft_mask")
        with torch.cuda.device(vimg.device):
            rc = lib.dibr_render_mesh_bwd(
                B, V, F, H, W, D, _ptr(g), _ptr(gm), _ptr(idx), _ptr(vimg), _ptr(faces32),
                _ptr(vfeat), shared, _ptr(mask), sigmainv, margin, knum, multiplier, eps,
                _ptr(grad_vimg), _ptr(grad_vfeat), _ptr(ws), ws.numel(), cap_r,
                cap_s, cap_p, _stream(vimg.device))
        _lib.check(rc, "dibr_render_mesh_bwd")
        return (None, None, None, grad_vimg, None, grad_vfeat, None, None, None, None, None, None)


def dibr_rasterization_mesh(height, width, vertices_camera, vertices_image, faces,
                            vertex_features, valid_faces=None, sigmainv=7000, boxlen=0.02,
                            knum=30, multiplier=None, eps=None, cuda_graph=False):
    r"""dibr_rasterization fed by the mesh itself (extension; not in the reference).

    Equivalent -- bit for bit in the forward -- to

        face_vertices_camera, face_vertices_image, face_normals = prepare_vertices(...)
        face_features = index_vertices_by_faces(vertex_features, faces)
        dibr_rasterization(height, width, face_vertices_camera[..., 2], face_vertices_image,
                           face_features, face_normals[..., 2], ...)

    (kaolin/render/mesh/utils.py:128-175, kaolin/render/mesh/dibr.py:119-209) without ever
    building the (batch_size, num_faces, 3, .) tensors: the kernels read the corners through
    ``faces`` and the backward accumulates vertex-level gradients.

    Args:
        height, width (int): size of the rendered images.
        vertices_camera (torch.FloatTensor): (batch_size, num_vertices, 3), camera space (the
            output of ``rotate_translate_points``); supplies depth and the normal test.
        vertices_image (torch.FloatTensor): (batch_size, num_vertices, 2), NDC in [-1, 1].
        faces (torch.IntTensor or LongTensor): (num_faces, 3).
        vertex_features (torch.FloatTensor or list): (batch_size, num_vertices, feature_dim),
            or (1, num_vertices, feature_dim) shared by all views (its gradient is then the
            sum over the views).
        valid_faces (torch.BoolTensor): (batch_size, num_faces); default: faces whose unit
            normal has z >= 0 in camera space, as dibr_rasterization uses.
        sigmainv, boxlen, knum, multiplier, eps: as dibr_rasterization.

    Returns:
        (torch.Tensor or tuple, torch.Tensor, torch.LongTensor): as dibr_rasterization.
    """
    _multiplier = 1000. if multiplier is None else multiplier
    _eps = 1e-8 if eps is None else eps
    _features = torch.cat(vertex_features, dim=-1) \
        if isinstance(vertex_features, (list, tuple)) else vertex_features
    if vertices_camera.dtype != torch.float32:
        raise TypeError("dibr_rasterization_mesh takes float32 tensors")
    faces32 = _prep_faces(faces, vertices_camera.shape[1])
    res = None
    if cuda_graph:
        from .dibr_graph import graphed_dibr_rasterization_mesh
        res = graphed_dibr_rasterization_mesh(
            height, width, vertices_camera, vertices_image, faces32, _features, valid_faces,
            (float(sigmainv), _margin(boxlen, _multiplier), int(knum), float(_multiplier),
             float(_eps)))
    if res is None:
        res = DibrRasterizationMeshB200.apply(
            height, width, vertices_camera, vertices_image, faces32, _features, valid_faces,
            sigmainv, boxlen, knum, _multiplier, _eps)
    image_features, soft_mask, face_idx = res
    if isinstance(vertex_features, (list, tuple)):
        _image_features = []
        cur_idx = 0
        for f in vertex_features:
            _image_features.append(image_features[..., cur_idx:cur_idx + f.shape[-1]])
            cur_idx += f.shape[-1]
        image_features = tuple(_image_features)
    return image_features, soft_mask, face_idx


class VertexTransformB200(Function):
    """The V-sized head of prepare_vertices (render/mesh/utils.py:158-170) as one kernel each
    way: camera transform + perspective projection.  Gradients reach the vertices only."""

    @staticmethod
    def forward(ctx, vertices, camera_proj, camera_rot, camera_trans, camera_transform, batch):
        v = _prep(vertices, "vertices")
        proj = _prep(camera_proj.reshape(-1), "camera_proj")
        rot = None if camera_rot is None else _prep(camera_rot, "camera_rot")
        trans = None if camera_trans is None else _prep(camera_trans.reshape(-1, 3), "camera_trans")
        tf = None if camera_transform is None else _prep(camera_transform, "camera_transform")
        if v.dim() != 3 or v.shape[-1] != 3 or v.shape[0] not in (1, batch) or proj.numel() != 3:
            raise RuntimeError("expected vertices (batch_size or 1, num_vertices, 3) and "
                               "camera_proj of 3 elements")
        if (rot is not None and rot.shape != (batch, 3, 3)) or \
                (trans is not None and trans.shape != (batch, 3)) or \
                (tf is not None and tf.shape != (batch, 4, 3)):
            raise RuntimeError("camera_rot (B,3,3), camera_trans (B,3) or camera_transform (B,4,3)")
        V = v.shape[1]
        shared = int(v.shape[0] == 1)
        dev = v.device
        vcam = torch.empty((batch, V, 3), dtype=torch.float32, device=dev)
        vimg = torch.empty((batch, V, 2), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().dibr_vertex_transform_fwd(
                batch, V, shared, _ptr(v), _ptr(rot), _ptr(trans), _ptr(tf), _ptr(proj),
                _ptr(vcam), _ptr(vimg), _stream(dev)), "dibr_vertex_transform_fwd")
        ctx.save_for_backward(vcam, proj, rot if rot is not None else tf)
        ctx.meta = (batch, V, shared, rot is not None, tuple(v.shape))
        ctx.set_materialize_grads(False)
        return vcam, vimg

    @staticmethod
    def backward(ctx, g_vcam, g_vimg):
        vcam, proj, mat = ctx.saved_tensors
        batch, V, shared, is_rot, vshape = ctx.meta
        if g_vcam is None and g_vimg is None:
            return (None,) * 6
        g_vcam = None if g_vcam is None else _prep(g_vcam, "grad_vertices_camera")
        g_vimg = None if g_vimg is None else _prep(g_vimg, "grad_vertices_image")
        gv = torch.empty(vshape, dtype=torch.float32, device=vcam.device)
        with torch.cuda.device(vcam.device):
            _lib.check(_lib.lib().dibr_vertex_transform_bwd(
                batch, V, shared, _ptr(vcam), _ptr(mat) if is_rot else None,
                None if is_rot else _ptr(mat), _ptr(proj), _ptr(g_vcam), _ptr(g_vimg), _ptr(gv),
                _stream(vcam.device)), "dibr_vertex_transform_bwd")
        return gv, None, None, None, None, None


def _camera_batch(camera_rot, camera_trans, camera_transform):
    if camera_transform is None:
        assert camera_trans is not None and camera_rot is not None, \
            "camera_transform or camera_trans and camera_rot must be defined"
        return camera_rot.shape[0]
    assert camera_trans is None and camera_rot is None, \
        "camera_trans and camera_rot must be None when camera_transform is defined"
    return camera_transform.shape[0]


def dibr_rasterization_vertices(height, width, vertices, faces, vertex_features, camera_proj,
                                camera_rot=None, camera_trans=None, camera_transform=None,
                                valid_faces=None, sigmainv=7000, boxlen=0.02, knum=30,
                                multiplier=None, eps=None, cuda_graph=False, exchange=None):
    r"""prepare_vertices + dibr_rasterization in one call (extension; not in the reference):
    world-space vertices and cameras in, images out, vertex-level gradients back.

    The camera arguments are prepare_vertices' (kaolin/render/mesh/utils.py:128-175); the
    camera transform and the projection run as one V-sized kernel, the renderer reads the
    corners through ``faces`` (see :func:`dibr_rasterization_mesh`).  ``vertices`` and
    ``vertex_features`` may have a leading dimension of 1: one mesh seen by all cameras, whose
    gradients are summed over the views.  Camera parameters receive no gradient (use
    ``prepare_vertices`` + ``dibr_rasterization_mesh`` for that).  ``cuda_graph=True`` replays
    the whole forward and the whole backward as one graph launch each.

    ``exchange`` (a ``parallel.VertexGradExchange``; multi-GPU, one process per GPU, each
    rendering its shard of the views of ONE shared mesh): the gradients of ``vertices`` and
    ``vertex_features`` -- both with a leading dimension of 1 -- come back already summed
    over all ranks, the path's only exchange.  With ``cuda_graph=True`` and peer-memory mode
    the backward kernels write straight into the exchange buffer and the all-reduce kernel is
    the last node of the backward graph: no copy, no extra launch.

    Returns:
        (torch.Tensor or tuple, torch.Tensor, torch.LongTensor): as dibr_rasterization.
    """
    batch = _camera_batch(camera_rot, camera_trans, camera_transform)
    _multiplier = 1000. if multiplier is None else multiplier
    _eps = 1e-8 if eps is None else eps
    _features = torch.cat(vertex_features, dim=-1) \
        if isinstance(vertex_features, (list, tuple)) else vertex_features
    if vertices.dtype != torch.float32:
        raise TypeError("dibr_rasterization_vertices takes float32 tensors")
    faces32 = _prep_faces(faces, vertices.shape[1])
    res = None
    fold = False
    if exchange is not None:
        if vertices.shape[0] != 1 or _features.shape[0] != 1:
            raise RuntimeError("exchange= needs ONE shared mesh: vertices (1, V, 3) and "
                               "vertex_features (1, V, D)")
        # every rank decides the same way (same shapes, same mode)
        pad = (vertices.numel() + 63) // 64 * 64
        fold = cuda_graph and exchange.prepare(pad + _features.numel(), vertices.device)
    if exchange is not None and not fold:
        # not folded into a graph: an identity node whose backward sums both gradients
        from ...parallel import SumGradsAcrossRanks
        vertices, _features = SumGradsAcrossRanks.apply(exchange, vertices, _features)
    if cuda_graph:
        from .dibr_graph import graphed_dibr_rasterization_vertices
        res = graphed_dibr_rasterization_vertices(
            height, width, vertices, faces32, _features, camera_proj, camera_rot, camera_trans,
            camera_transform, valid_faces, batch,
            (float(sigmainv), _margin(boxlen, _multiplier), int(knum), float(_multiplier),
             float(_eps)), exchange if fold else None)
    if res is None and fold:
        # first sighting of a problem whose graph will hold the exchange: eager this once
        from ...parallel import SumGradsAcrossRanks
        vertices, _features = SumGradsAcrossRanks.apply(exchange, vertices, _features)
    if res is None:
        vcam, vimg = VertexTransformB200.apply(vertices, camera_proj, camera_rot, camera_trans,
                                               camera_transform, batch)
        res = DibrRasterizationMeshB200.apply(
            height, width, vcam, vimg, faces32, _features, valid_faces, sigmainv, boxlen, knum,
            _multiplier, _eps)
    image_features, soft_mask, face_idx = res
    if isinstance(vertex_features, (list, tuple)):
        _image_features = []
        cur_idx = 0
        for f in vertex_features:
            _image_features.append(image_features[..., cur_idx:cur_idx + f.shape[-1]])
            cur_idx += f.shape[-1]
        image_features = tuple(_image_features)
    return image_features, soft_mask, face_idx
