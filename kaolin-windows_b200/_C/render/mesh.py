"""kaolin._C.render.mesh.{packed_rasterize_forward_cuda, rasterize_backward_cuda,
dibr_soft_mask_forward_cuda, dibr_soft_mask_backward_cuda} with the reference's exact
signatures, output allocation and return order (rasterization.cpp:49-168,
dibr_soft_mask.cpp:48-183), over the C ABI of libdibr_b200."""
import torch

from ... import _lib, _workspace
from ...render.mesh.rasterization import _ptr, _stream

__all__ = ["packed_rasterize_forward_cuda", "rasterize_backward_cuda",
           "dibr_soft_mask_forward_cuda", "dibr_soft_mask_backward_cuda",
           "deftet_sparse_render_forward_cuda", "deftet_sparse_render_backward_cuda"]


def _check_all(fn, named, dtype=None):
    """at::checkAllSameGPU + at::checkAllContiguous (rasterization.cpp:70-77)."""
    dev = None
    for name, t in named:
        if not t.is_cuda:
            raise RuntimeError("%s: expected %s to be a CUDA tensor (no CPU path)" % (fn, name))
        if dev is None:
            dev = t.device
        elif t.device != dev:
            raise RuntimeError("%s: expected all tensors on the same GPU, %s is on %s"
                               % (fn, name, t.device))
        if not t.is_contiguous():
            raise RuntimeError("%s: expected %s to be contiguous" % (fn, name))
    return dev


def _check_size(fn, name, t, size):
    if tuple(t.shape) != tuple(size):
        raise RuntimeError("%s: expected %s of size %s, got %s"
                           % (fn, name, tuple(size), tuple(t.shape)))


def _check_float(fn, *ts):
    """AT_DISPATCH_FLOATING_TYPES (rasterization_cuda.cu:30-39): float or double, all alike.
    Returns True for double."""
    dt = ts[0].dtype
    for t in ts:
        if t.dtype not in (torch.float32, torch.float64):
            raise RuntimeError("%s not implemented for '%s'" % (fn, t.dtype))
        if t.dtype != dt:
            raise RuntimeError("%s: expected all floating tensors of dtype %s, got %s"
                               % (fn, dt, t.dtype))
    return dt == torch.float64


def _aligned(t):
    return t if t.data_ptr() % 16 == 0 else t.clone()


def packed_rasterize_forward_cuda(height, width, face_vertices_z, face_vertices_image,
                                  face_bboxes, face_features, first_idx_face_per_mesh,
                                  multiplier, eps):
    fn = "packed_rasterize_forward_cuda"
    dev = _check_all(fn, [("face_vertices_z", face_vertices_z),
                          ("face_vertices_image", face_vertices_image),
                          ("face_bboxes", face_bboxes), ("face_features", face_features),
                          ("first_idx_face_per_mesh", first_idx_face_per_mesh)])
    f64 = _check_float(fn, face_vertices_z, face_vertices_image, face_bboxes, face_features)
    num_faces = face_vertices_z.shape[0]
    batch_size = first_idx_face_per_mesh.shape[0] - 1
    feat_dim = face_features.shape[2]
    _check_size(fn, "face_vertices_z", face_vertices_z, (num_faces, 3))
    _check_size(fn, "face_vertices_image", face_vertices_image, (num_faces, 3, 2))
    _check_size(fn, "face_bboxes", face_bboxes, (num_faces, 4))
    _check_size(fn, "face_features", face_features, (num_faces, 3, feat_dim))
    _check_size(fn, "first_idx_face_per_mesh", first_idx_face_per_mesh, (batch_size + 1,))
    first = first_idx_face_per_mesh.to(torch.long)
    z, img, bbox, feat = (_aligned(t) for t in (face_vertices_z, face_vertices_image,
                                                  face_bboxes, face_features))
    idx = torch.empty((batch_size, height, width), dtype=torch.long, device=dev)
    weights = torch.empty((batch_size, height, width, 3), dtype=z.dtype, device=dev)
    interp = torch.empty((batch_size, height, width, feat_dim), dtype=z.dtype, device=dev)
    lib = _lib.lib()
    with torch.cuda.device(dev):
        st = _stream(dev)
        if f64:
            img32, bbox32 = img.float(), bbox.float()    # binning only; the doubles decide

            def call64(ws, ws_bytes, cap_r, cap_s, cap_p, needed):
                return lib.dibr_packed_rasterize_fwd_f64(
                    batch_size, num_faces, int(height), int(width), feat_dim, _ptr(z), _ptr(img),
                    _ptr(bbox), _ptr(feat), _ptr(img32), _ptr(bbox32), _ptr(first),
                    float(multiplier), float(eps), _ptr(idx), _ptr(interp), _ptr(weights),
                    ws, ws_bytes, cap_r, needed, st)

            _workspace.run("packed64", dev, batch_size, num_faces, int(height), int(width), True,
                           False, call64)
            return [interp, idx, weights]

        def call(ws, ws_bytes, cap_r, cap_s, cap_p, needed):
            return lib.dibr_rasterize_fwd(
                batch_size, num_faces, int(height), int(width), feat_dim, _ptr(z), _ptr(img),
                _ptr(feat), None, None, 1, _ptr(bbox), _ptr(first), float(multiplier), float(eps),
                _ptr(idx), _ptr(interp), _ptr(weights), ws, ws_bytes, cap_r, needed, st)

        _workspace.run("packed", dev, batch_size, num_faces, int(height), int(width), True, False,
                       call)
    return [interp, idx, weights]


def rasterize_backward_cuda(grad_interpolated_features, interpolated_features,
                            selected_face_idx, output_weights, face_vertices_image,
                            face_features, eps):
    fn = "rasterize_backward_cuda"
    dev = _check_all(fn, [("grad_interpolated_features", grad_interpolated_features),
                          ("interpolated_features", interpolated_features),
                          ("selected_face_idx", selected_face_idx),
                          ("output_weights", output_weights),
                          ("face_vertices_image", face_vertices_image),
                          ("face_features", face_features)])
    f64 = _check_float(fn, grad_interpolated_features, output_weights, face_vertices_image,
                       face_features)
    B, H, W, D = grad_interpolated_features.shape
    F = face_vertices_image.shape[1]
    _check_size(fn, "interpolated_features", interpolated_features, (B, H, W, D))
    _check_size(fn, "selected_face_idx", selected_face_idx, (B, H, W))
    _check_size(fn, "output_weights", output_weights, (B, H, W, 3))
    _check_size(fn, "face_vertices_image", face_vertices_image, (B, F, 3, 2))
    _check_size(fn, "face_features", face_features, (B, F, 3, D))
    img = _aligned(face_vertices_image)
    grad_img = torch.zeros_like(img)
    grad_feat = torch.zeros_like(face_features)
    with torch.cuda.device(dev):
        if f64:
            rc = _lib.lib().dibr_rasterize_bwd_f64(
                B, F, H, W, D, _ptr(grad_interpolated_features), _ptr(selected_face_idx),
                _ptr(output_weights), _ptr(img), _ptr(face_features), float(eps),
                _ptr(grad_img), _ptr(grad_feat), _stream(dev))
            _lib.check(rc, fn)
            return [grad_img, grad_feat]
        rc = _lib.lib().dibr_rasterize_bwd(
            B, F, H, W, D, _ptr(grad_interpolated_features), _ptr(selected_face_idx),
            _ptr(output_weights), _ptr(img), _ptr(face_features), 1.0, float(eps),
            _ptr(grad_img), _ptr(grad_feat), _stream(dev))
    _lib.check(rc, fn)
    return [grad_img, grad_feat]


def dibr_soft_mask_forward_cuda(face_vertices_image, face_large_bboxes, selected_face_idx,
                                sigmainv, knum, multiplier):
    fn = "dibr_soft_mask_forward_cuda"
    dev = _check_all(fn, [("face_vertices_image", face_vertices_image),
                          ("face_bboxes", face_large_bboxes),
                          ("selected_face_idx", selected_face_idx)])
    f64 = _check_float(fn, face_vertices_image, face_large_bboxes)
    B, F = face_vertices_image.shape[:2]
    H, W = selected_face_idx.shape[1:3]
    _check_size(fn, "face_vertices_image", face_vertices_image, (B, F, 3, 2))
    _check_size(fn, "face_bboxes", face_large_bboxes, (B, F, 4))
    _check_size(fn, "selected_face_idx", selected_face_idx, (B, H, W))
    img, bbox = _aligned(face_vertices_image), _aligned(face_large_bboxes)
    knum = int(knum)
    cidx = torch.full((B, H, W, knum), -1, dtype=torch.long, device=dev)
    prob = torch.zeros((B, H, W, knum), dtype=img.dtype, device=dev)
    typ = torch.zeros((B, H, W, knum), dtype=torch.uint8, device=dev)
    mask = torch.zeros((B, H, W), dtype=img.dtype, device=dev)
    lib = _lib.lib()
    with torch.cuda.device(dev):
        st = _stream(dev)
        if f64:
            img32, bbox32 = img.float(), bbox.float()    # binning only; the doubles decide

            def call64(ws, ws_bytes, cap_r, cap_s, cap_p, needed):
                return lib.dibr_softmask_fwd_f64(
                    B, F, H, W, _ptr(img), _ptr(bbox), _ptr(img32), _ptr(bbox32),
                    _ptr(selected_face_idx), float(sigmainv), knum, float(multiplier), _ptr(mask),
                    _ptr(prob), _ptr(cidx), _ptr(typ), ws, ws_bytes, cap_s, needed, st)

            _workspace.run("soft64", dev, B, F, H, W, False, True, call64)
            return [mask, prob, cidx, typ]

        def call(ws, ws_bytes, cap_r, cap_s, cap_p, needed):
            return lib.dibr_softmask_fwd(
                B, F, H, W, _ptr(img), 1, _ptr(bbox), _ptr(selected_face_idx), float(sigmainv),
                0.0, knum, float(multiplier), _ptr(mask), _ptr(prob), _ptr(cidx), _ptr(typ),
                ws, ws_bytes, cap_r, cap_s, 0, needed, st)

        _workspace.run("soft_dense", dev, B, F, H, W, False, True, call)
    return [mask, prob, cidx, typ]


def dibr_soft_mask_backward_cuda(grad_soft_mask, soft_mask, selected_face_idx, close_face_prob,
                                 close_face_idx, close_face_dist_type, face_vertices_image,
                                 sigmainv, multiplier):
    fn = "dibr_soft_mask_backward_cuda"
    dev = _check_all(fn, [("grad_soft_mask", grad_soft_mask), ("soft_mask", soft_mask),
                          ("selected_face_idx", selected_face_idx),
                          ("close_face_prob", close_face_prob),
                          ("close_face_idx", close_face_idx),
                          ("close_face_dist_type", close_face_dist_type),
                          ("face_vertices_image", face_vertices_image)])
    f64 = _check_float(fn, grad_soft_mask, soft_mask, close_face_prob, face_vertices_image)
    B, F = face_vertices_image.shape[:2]
    H, W = selected_face_idx.shape[1:3]
    K = close_face_idx.shape[-1]
    _check_size(fn, "grad_soft_mask", grad_soft_mask, (B, H, W))
    _check_size(fn, "soft_mask", soft_mask, (B, H, W))
    _check_size(fn, "selected_face_idx", selected_face_idx, (B, H, W))
    _check_size(fn, "close_face_prob", close_face_prob, (B, H, W, K))
    _check_size(fn, "close_face_idx", close_face_idx, (B, H, W, K))
    _check_size(fn, "close_face_dist_type", close_face_dist_type, (B, H, W, K))
    _check_size(fn, "face_vertices_image", face_vertices_image, (B, F, 3, 2))
    grad_img = torch.zeros_like(face_vertices_image)
    with torch.cuda.device(dev):
        fnc = _lib.lib().dibr_softmask_bwd_dense_f64 if f64 else _lib.lib().dibr_softmask_bwd_dense
        rc = fnc(
            B, F, H, W, K, _ptr(grad_soft_mask), _ptr(soft_mask), _ptr(selected_face_idx),
            _ptr(close_face_prob), _ptr(close_face_idx), _ptr(close_face_dist_type),
            _ptr(face_vertices_image), float(sigmainv), float(multiplier), _ptr(grad_img),
            _stream(dev))
    _lib.check(rc, fn)
    return grad_img


def deftet_sparse_render_forward_cuda(face_vertices_z, face_vertices_image, face_bboxes,
                                      pixel_coords, pixel_depth_ranges, knum, eps):
    """deftet.cpp:47-107: the raw op -- unsorted slots in ascending face order."""
    fn = "deftet_sparse_render_forward_cuda"
    dev = _check_all(fn, [("face_vertices_z", face_vertices_z),
                          ("face_vertices_image", face_vertices_image),
                          ("face_bboxes", face_bboxes), ("pixel_coords", pixel_coords),
                          ("pixel_depth_ranges", pixel_depth_ranges)])
    f64 = _check_float(fn, face_vertices_z, face_vertices_image, face_bboxes, pixel_coords,
                       pixel_depth_ranges)
    B, F = face_vertices_z.shape[:2]
    P = pixel_coords.shape[1]
    knum = int(knum)
    _check_size(fn, "face_vertices_z", face_vertices_z, (B, F, 3))
    _check_size(fn, "face_vertices_image", face_vertices_image, (B, F, 3, 2))
    _check_size(fn, "face_bboxes", face_bboxes, (B, F, 4))
    _check_size(fn, "pixel_coords", pixel_coords, (B, P, 2))
    _check_size(fn, "pixel_depth_ranges", pixel_depth_ranges, (B, P, 2))
    z, img, bbox, pix, rng = (_aligned(t) for t in (face_vertices_z, face_vertices_image,
                                                     face_bboxes, pixel_coords,
                                                     pixel_depth_ranges))
    idx = torch.empty((B, P, knum), dtype=torch.long, device=dev)
    depth = torch.empty((B, P, knum), dtype=z.dtype, device=dev)
    w0 = torch.empty((B, P, knum), dtype=z.dtype, device=dev)
    w1 = torch.empty((B, P, knum), dtype=z.dtype, device=dev)
    from ...render.mesh.deftet import workspace
    with torch.cuda.device(dev):
        ws, ws_bytes = workspace(dev, B, F, P)
        rc = _lib.lib().dibr_deftet_fwd_raw(
            1 if f64 else 0, B, F, P, knum, _ptr(z), _ptr(img), _ptr(bbox), _ptr(pix), _ptr(rng),
            float(eps), _ptr(idx), _ptr(depth), _ptr(w0), _ptr(w1), _ptr(ws), ws_bytes,
            _stream(dev))
    _lib.check(rc, fn)
    return [idx, depth, w0, w1]


def deftet_sparse_render_backward_cuda(grad_interpolated_features, face_idx, weights,
                                       face_vertices_image, face_features, eps):
    """deftet.cpp:109-165."""
    fn = "deftet_sparse_render_backward_cuda"
    dev = _check_all(fn, [("grad_interpolated_features", grad_interpolated_features),
                          ("face_idx", face_idx), ("weights", weights),
                          ("face_vertices_image", face_vertices_image),
                          ("face_features", face_features)])
    f64 = _check_float(fn, grad_interpolated_features, weights, face_vertices_image,
                       face_features)
    B, P, K, D = grad_interpolated_features.shape
    F = face_vertices_image.shape[1]
    _check_size(fn, "face_idx", face_idx, (B, P, K))
    _check_size(fn, "weights", weights, (B, P, K, 3))
    _check_size(fn, "face_vertices_image", face_vertices_image, (B, F, 3, 2))
    _check_size(fn, "face_features", face_features, (B, F, 3, D))
    img = _aligned(face_vertices_image)
    grad_img = torch.empty_like(img)
    grad_feat = torch.empty_like(face_features)
    with torch.cuda.device(dev):
        rc = _lib.lib().dibr_deftet_bwd(
            1 if f64 else 0, B, F, P, K, D, _ptr(grad_interpolated_features), _ptr(face_idx),
            _ptr(weights), _ptr(img), _ptr(face_features), float(eps), _ptr(grad_img),
            _ptr(grad_feat), _stream(dev))
    _lib.check(rc, fn)
    return [grad_img, grad_feat]
