"""Known-answer vectors of the reference's own tests for texture_mapping
(tests/python/kaolin/render/mesh/test_utils.py:24-106) and mask_iou
(tests/python/kaolin/metrics/test_render.py:25-53), restated as data."""
import torch


def texture_maps():
    l1 = torch.tensor([[11., 12., 13., 14., 15.], [21., 22., 23., 24., 25.],
                       [31., 32., 33., 34., 35.], [41., 42., 43., 44., 45.]])
    t1 = torch.stack((l1, l1 + 100)).unsqueeze(1)            # (2, 1, 4, 5)
    return t1, torch.cat((t1, -t1, t1), dim=1)               # 1 and 3 channels


def sparse_coords():
    c = torch.tensor([[0.0, 0.0], [1.0, 1.0], [0, 1.0], [0.5, 0.5]])
    return torch.stack((c, torch.flip(c, dims=(0,))))         # (2, 4, 2)


def dense_coords():
    c = torch.tensor([[[0.0, 0.0], [0.25, 0.0], [0.5, 0.0], [0.75, 0.0], [1.0, 0.0]],
                      [[0.0, 1 / 8], [0.25, 1 / 8], [0.5, 1 / 8], [0.75, 1 / 8], [1.0, 1 / 8]]])
    return torch.stack((c, torch.flip(c, dims=(0,))))         # (2, 2, 5, 2)


def sparse_expected(mode, channels):
    if mode == 'nearest':
        e = torch.tensor([[41., 15, 11, 33], [133, 111, 115, 141]])
    else:
        e = torch.tensor([[41., 15, 11, 28], [128, 111, 115, 141]])
    return e.unsqueeze(-1) if channels == 1 else torch.stack([e, -e, e], dim=-1)


def dense_expected(mode):
    base = torch.tensor([41., 42., 43., 44., 45.]) if mode == 'nearest' else \
        torch.tensor([41., 41.75, 43., 44.25, 45.])
    e = torch.stack([base, base + 100], dim=0)
    return torch.stack([e, -e, e], dim=-1).reshape(2, 1, -1, 3).repeat(1, 2, 1, 1)


def iou_masks():
    lhs = torch.tensor([[[0., 0.2, 0.1, 1.], [0.5, 0.5, 0.9, 0.9], [0., 1., 1., 0.9],
                         [0.8, 0.7, 0.2, 0.1]],
                        [[1., 1., 1., 1.], [1., 1., 1., 1.], [1., 1., 1., 1.], [1., 1., 1., 1.]]])
    rhs = torch.tensor([[[0.1, 0.3, 0.3, 0.9], [0.5, 0.5, 1., 0.3], [0., 0.9, 0.9, 0.8],
                         [1., 1., 0., 0.]],
                        [[0.3, 0.6, 0.7, 0.7], [0.8, 0.9, 0.9, 1.], [1., 0.9, 0.9, 0.5],
                         [0.8, 0.7, 0.8, 0.5]]])
    return lhs, rhs          # expected loss 0.3105 (allclose)
