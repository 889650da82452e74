"""The epilogue oracle (oracle/mesh_oracle.py: texture_mapping, mask_iou) against the
known-answer vectors of the reference's own tests -- CPU only."""
import pytest
import torch

from oracle import mesh_oracle as mo
import epilogue_kats as kats


@pytest.mark.parametrize("mode", ["nearest", "bilinear"])
def test_oracle_texture_mapping_kats(mode):
    t1, t3 = kats.texture_maps()
    assert torch.equal(mo.texture_mapping(kats.sparse_coords(), t1, mode), kats.sparse_expected(mode, 1))
    assert torch.equal(mo.texture_mapping(kats.sparse_coords(), t3, mode), kats.sparse_expected(mode, 3))
    assert torch.equal(mo.texture_mapping(kats.dense_coords(), t3, mode), kats.dense_expected(mode))


def test_oracle_mask_iou_kat():
    lhs, rhs = kats.iou_masks()
    assert torch.allclose(mo.mask_iou(lhs, rhs), torch.tensor([0.3105]))
