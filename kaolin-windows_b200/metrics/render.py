"""kaolin.metrics.render.mask_iou on B200 (kaolin/metrics/render.py:18-40): the loss the
DIB-R loop puts on the soft mask.  Both per-view sums come from ONE pass over the two
masks (the reference materialises their product and sum and reduces each), the backward is
one elementwise kernel; the B-sized tail (ratio, mean) is a one-warp kernel."""
import ctypes

import torch
from torch.autograd import Function

from .. import _lib

__all__ = ["mask_iou"]


def _ptr(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _stream(device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


class _MaskIoU(Function):
    @staticmethod
    def forward(ctx, lhs, rhs):
        B = lhs.shape[0]
        npix = lhs.numel() // B if B > 0 else 0
        sums = torch.empty((B, 2), dtype=torch.float32, device=lhs.device)
        if B == 0:                       # the reference's mean over no views
            ctx.save_for_backward(lhs, rhs, sums)
            return torch.full((), float("nan"), dtype=torch.float32, device=lhs.device)
        with torch.cuda.device(lhs.device):
            _lib.check(_lib.lib().dibr_mask_iou_sums(B, npix, _ptr(lhs), _ptr(rhs), _ptr(sums),
                                                     _stream(lhs.device)), "mask_iou")
            loss = torch.empty((), dtype=torch.float32, device=lhs.device)
            _lib.check(_lib.lib().dibr_mask_iou_loss(B, _ptr(sums), _ptr(loss),
                                                     _stream(lhs.device)), "mask_iou")  # render.py:37-40
        ctx.save_for_backward(lhs, rhs, sums)
        return loss

    @staticmethod
    def backward(ctx, grad_loss):
        lhs, rhs, sums = ctx.saved_tensors
        B = lhs.shape[0]
        npix = lhs.numel() // B if B > 0 else 0
        g_l = torch.empty_like(lhs) if ctx.needs_input_grad[0] else None
        g_r = torch.empty_like(rhs) if ctx.needs_input_grad[1] else None
        g = grad_loss.to(torch.float32).contiguous()
        with torch.cuda.device(lhs.device):
            _lib.check(_lib.lib().dibr_mask_iou_bwd(B, npix, _ptr(lhs), _ptr(rhs), _ptr(sums),
                                                    _ptr(g), _ptr(g_l), _ptr(g_r),
                                                    _stream(lhs.device)), "mask_iou backward")
        return g_l, g_r


def mask_iou(lhs_mask, rhs_mask):
    r"""Compute the Intersection over Union of two segmentation masks.

    Args:
        lhs_mask (torch.FloatTensor): of shape (batch_size, height, width).
        rhs_mask (torch.FloatTensor): of shape (batch_size, height, width).

    Returns:
        (torch.FloatTensor): The IoU loss, as a torch scalar.
    """
    batch_size, height, width = lhs_mask.shape
    assert rhs_mask.shape == lhs_mask.shape
    for t, name in ((lhs_mask, "lhs_mask"), (rhs_mask, "rhs_mask")):
        if not t.is_cuda:
            raise RuntimeError("%s must be a CUDA tensor (there is no CPU path)" % name)
        if t.dtype != torch.float32:
            raise TypeError("%s must be float32, got %s" % (name, t.dtype))
    return _MaskIoU.apply(lhs_mask.contiguous(), rhs_mask.contiguous())
