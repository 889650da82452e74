"""The peer-memory exchange (csrc/dibr_peer.cu via parallel.MeshGradExchange) against NCCL
on two GPUs: values, bit-identity across ranks, eager / captured / replayed calls.  Needs a
box with at least two GPUs; skipped otherwise (tools/check_peer_exchange.py is the body)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_peer_exchange_matches_nccl():
    r = subprocess.run(
        [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
         "--master-addr", "127.0.0.1", "--master-port", "29533",
         os.path.join(ROOT, "tools", "check_peer_exchange.py")],
        cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and '"ok": true' in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
