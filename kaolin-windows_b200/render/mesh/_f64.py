"""float64 inputs: the reference dispatches float and double to the same kernels
(rasterization_cuda.cu:30-39, dibr_soft_mask_cuda.cu:211,382) and tests both.  Here double
takes the reference's own route -- its torch-side preparation (RasterizeCuda.forward,
rasterization.py:289-369; DibrSoftMaskCuda.forward, dibr.py:29-55) around the four raw ops
-- with the raw ops served by the double kernels of csrc/dibr_f64.cu through the
`_C.render.mesh` shim.  Built for the reference's test-sized double workloads, not for
throughput (per-pixel dense soft-mask state as in the reference).
"""
import torch
from torch.autograd import Function

from ... import _C

__all__ = ["RasterizeF64", "DibrSoftMaskF64"]


def _need_cuda(t, name):
    if not t.is_cuda:
        raise RuntimeError("%s must be a CUDA tensor (there is no CPU path)" % name)


class RasterizeF64(Function):
    """RasterizeCuda for double tensors (rasterization.py:260-388)."""

    @staticmethod
    def forward(ctx, height, width, face_vertices_z, face_vertices_image, face_features,
                valid_faces, multiplier, eps):
        _need_cuda(face_vertices_z, "face_vertices_z")
        B, F = face_vertices_z.shape[:2]
        D = face_features.shape[-1]
        dev = face_vertices_z.device
        face_features = face_features.contiguous()
        face_vertices_image = face_vertices_image.contiguous()
        if valid_faces is None:        # every face, in order (:309-323)
            face_of = torch.arange(F, dtype=torch.long, device=dev).repeat(B)
            img = face_vertices_image.reshape(B * F, 3, 2)
            z = face_vertices_z.reshape(B * F, 3)
            feat = face_features.reshape(B * F, 3, D)
            per_mesh = torch.full((B,), F, dtype=torch.long, device=dev)
        else:                          # valid faces packed mesh after mesh (:325-332)
            if valid_faces.shape != (B, F):
                raise RuntimeError("valid_faces must be of shape (batch_size, num_faces)")
            mesh_of, face_of = torch.where(valid_faces)
            img = face_vertices_image[mesh_of, face_of]
            z = face_vertices_z[mesh_of, face_of]
            feat = face_features[mesh_of, face_of]
            per_mesh = valid_faces.reshape(B, -1).sum(dim=1)
        first = torch.zeros(B + 1, dtype=torch.long, device=dev)
        torch.cumsum(per_mesh, dim=0, out=first[1:])
        img = img * multiplier                                             # :337
        bboxes = torch.cat((img.min(dim=1)[0], img.max(dim=1)[0]), dim=1)  # :342-344
        interp, packed_idx, weights = _C.render.mesh.packed_rasterize_forward_cuda(
            int(height), int(width), z.contiguous(), img.contiguous(), bboxes.contiguous(),
            feat.contiguous(), first, multiplier, eps)
        # packed index -> index within the mesh, background stays -1 (:357-363)
        hit = packed_idx >= 0
        if face_of.numel() > 0:
            face_idx = face_of[((packed_idx + first[:-1].reshape(-1, 1, 1)) * hit).reshape(-1)]
            face_idx = face_idx.reshape(packed_idx.shape).contiguous()
            face_idx[~hit] = -1
        else:
            face_idx = packed_idx
        ctx.save_for_backward(interp, face_idx, weights, face_vertices_image, face_features)
        ctx.mark_non_differentiable(face_idx)
        ctx.eps = eps
        return interp, face_idx

    @staticmethod
    def backward(ctx, grad_interpolated_features, grad_face_idx):
        interp, face_idx, weights, face_vertices_image, face_features = ctx.saved_tensors
        grad_img, grad_feat = _C.render.mesh.rasterize_backward_cuda(
            grad_interpolated_features.contiguous(), interp, face_idx, weights,
            face_vertices_image, face_features, ctx.eps)
        return None, None, None, grad_img, grad_feat, None, None, None


class DibrSoftMaskF64(Function):
    """DibrSoftMaskCuda for double tensors (dibr.py:27-73)."""

    @staticmethod
    def forward(ctx, face_vertices_image, selected_face_idx, sigmainv, boxlen, knum, multiplier):
        _need_cuda(face_vertices_image, "face_vertices_image")
        img = face_vertices_image.contiguous() * multiplier                # :31-32
        sel = selected_face_idx.contiguous()
        lo = img.min(dim=-2)[0] - boxlen * multiplier                      # :34-39
        hi = img.max(dim=-2)[0] + boxlen * multiplier
        mask, prob, cidx, typ = _C.render.mesh.dibr_soft_mask_forward_cuda(
            img, torch.cat([lo, hi], dim=-1).contiguous(), sel, sigmainv, knum, multiplier)
        ctx.multiplier, ctx.sigmainv = multiplier, sigmainv
        ctx.save_for_backward(mask, img, sel, prob, cidx, typ)
        return mask

    @staticmethod
    def backward(ctx, grad_soft_mask):
        mask, img, sel, prob, cidx, typ = ctx.saved_tensors
        grad_img = _C.render.mesh.dibr_soft_mask_backward_cuda(
            grad_soft_mask.contiguous(), mask, sel, prob, cidx, typ, img, ctx.sigmainv,
            ctx.multiplier)
        return grad_img, None, None, None, None, None
