"""The deftet oracle (oracle/deftet_oracle.{cpp,py}) against (a) the known answers of the
reference's own tests and (b) outputs of the reference's pure-torch
_naive_deftet_sparse_render generated from /root/reference (tests/golden/deftet_naive.npz,
script tests/golden/make_golden.py) -- CPU only."""
import os

import numpy as np
import pytest
import torch

from oracle import deftet_oracle as do
from workloads import fixture_geometry
import deftet_kats as kats

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
@pytest.mark.parametrize("case,lo", [("full", -4.), ("restricted", -2.1)])
def test_oracle_reference_kats(dtype, case, lo):
    pix, z, img, feats = kats.inputs(dtype)
    out, idx, _ = do.deftet_sparse_render_forward(pix, kats.ranges(lo, dtype), z, img,
                                                  torch.cat(feats, dim=-1), 5)
    gt_idx, gt_f0, gt_f1 = kats.expected(case, dtype)
    assert torch.equal(idx, gt_idx)
    assert torch.allclose(out[..., :1], gt_f0)
    assert torch.allclose(out[..., 1:], gt_f1, atol=3e-3, rtol=1e-5)   # test_deftet.py:161-165


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
@pytest.mark.parametrize("centre", [0, 1])
def test_oracle_vs_naive_golden(dtype, centre):
    g = np.load(os.path.join(GOLDEN, "deftet_naive.npz"))
    tag = str(dtype)[-7:]
    key = "%s_centre%d_k20" % (tag, centre)
    geo = fixture_geometry(batch_size=3, dtype=dtype)
    z, img, uvs = geo["face_vertices_z"], geo["face_vertices_image"], geo["face_uvs"]
    pix = torch.from_numpy(g[tag + "_pixel_coords"])
    rr = torch.from_numpy(g[key + "_ranges"])
    out, idx, weights = do.deftet_sparse_render_forward(pix, rr, z, img, uvs, 20)
    assert torch.equal(idx, torch.from_numpy(g[key + "_idx"]).long())
    assert torch.allclose(out, torch.from_numpy(g[key + "_feat"]), rtol=1e-4, atol=1e-4)
    gi, gf = do.deftet_sparse_render_backward(torch.from_numpy(g[key + "_grad_out"]), idx,
                                              weights, img, uvs)
    # bars of the reference's own test (test_deftet.py:482-487)
    assert torch.allclose(gi, torch.from_numpy(g[key + "_grad_img"]), rtol=5e-3, atol=5e-3)
    assert torch.allclose(gf, torch.from_numpy(g[key + "_grad_feat"]), rtol=1e-3, atol=1e-3)


def test_oracle_keeps_first_knum_in_face_order():
    """deftet_cuda.cu:163-187: beyond knum hits the FIRST knum faces (ascending index) stay,
    then they are ordered by depth."""
    pix, z, img, feats = kats.inputs()
    out, idx, _ = do.deftet_sparse_render_forward(pix, kats.ranges(-4.), z, img, feats[1], 2)
    assert idx[1, 2].tolist() == [1, 0]      # faces 0, 1 kept (2 is nearest but third)
    assert idx[1, 0].tolist() == [0, 1]
