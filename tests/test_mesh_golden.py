"""The mesh-side oracle (oracle/mesh_oracle.py) against the golden vectors generated from
the reference's own prepare_vertices (tests/golden/make_golden.py) -- CPU only."""
import os

import numpy as np
import pytest
import torch

from oracle import mesh_oracle as mo

GOLD = os.path.join(os.path.dirname(__file__), "golden", "prepare_vertices.npz")


def load():
    d = np.load(GOLD)
    return {k: torch.from_numpy(d[k]) for k in d.files}


def run(fn, d, mode, device="cpu"):
    leaves = {"vertices": d["vertices"], "camera_proj": d["camera_proj"]}
    if mode == "rot_trans":
        leaves.update(camera_rot=d["camera_rot"], camera_trans=d["camera_trans"])
    else:
        leaves.update(camera_transform=d["camera_transform"])
    leaves = {k: v.to(device).clone().requires_grad_(True) for k, v in leaves.items()}
    kw = {k: v for k, v in leaves.items() if k not in ("vertices", "camera_proj")}
    res = fn(leaves["vertices"], d["faces"].long().to(device), leaves["camera_proj"], **kw)
    cot = [d["cot%d" % i].to(device) for i in range(3)]
    torch.autograd.backward(res, cot)
    return res, {k: v.grad for k, v in leaves.items()}


def compare(res, grads, d, mode, rtol):
    for name, o in zip(("fvc", "fvi", "fn"), res):
        want = d["%s_%s" % (mode, name)]
        assert torch.allclose(o.detach().cpu(), want, rtol=rtol, atol=rtol), name
    for k, g in grads.items():
        want = d["%s_grad_%s" % (mode, k)]
        scale = float(want.abs().max())
        assert float((g.cpu() - want).abs().max()) <= rtol * 10 * scale, k


@pytest.mark.parametrize("mode", ["rot_trans", "transform"])
def test_oracle_matches_reference_prepare_vertices(mode):
    d = load()
    res, grads = run(mo.prepare_vertices, d, mode)
    compare(res, grads, d, mode, 1e-6)
