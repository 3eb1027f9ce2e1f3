"""2-rank NCCL check on a box with >= 2 GPUs (skipped otherwise): the path-sharded update
equals the single-GPU update within the same tolerances (rank-count invariance)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("peer", ["1", "0"])
def test_two_gpu_update_matches_one_gpu(peer):
    """peer=1: the exchanges run as csrc/peer.cu kernels over NVLink-mapped blocks (required, no silent
    NCCL fallback); peer=0: the same update over NCCL all-reduces."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", "2973" + peer, os.path.join(ROOT, "tests", "multi_gpu_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=dict(os.environ, MJRL_B200_PEER=peer))
    assert "MULTI_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
    assert ("peer=[active" in r.stdout) == (peer == "1"), r.stdout[-500:]
    assert "TRPO_DAPG_OK" in r.stdout and "PPO_OK" in r.stdout, r.stdout[-1500:]
    print("\n".join(r.stdout.strip().splitlines()[-3:]))
