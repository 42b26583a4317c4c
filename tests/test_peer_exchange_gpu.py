"""View-parallel gradient sums over peer memory (PeerGradientExchange: ggs_rasterize_backward_peer +
ggs_view_sum_peer) against the same step with NCCL collectives, two ranks on one node.  Needs two GPUs: skipped on a
single-GPU box (tools/peer_check.py is the same check, run under torchrun at N = 2 and N = 8 during development:
profiles/r01_peer_check_n8_c3.log)."""
import os
import subprocess
import sys

import pytest
import torch

import helpers as H

pytestmark = pytest.mark.gpu


def test_peer_exchange_matches_nccl_two_ranks():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs on one node")
    env = dict(os.environ, PEER_WORKLOAD="c1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29533", os.path.join(H.ROOT, "tools", "peer_check.py")]
    out = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert "OK" in out.stdout and "MISMATCH" not in out.stdout
