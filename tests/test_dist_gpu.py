"""Multi-GPU parity of the PPO update (needs two visible GPUs, skipped otherwise): the minibatch
sharded over two ranks -- NCCL all-reduce of the flat gradient and of the global denominators -- must give the losses and
gradient of the reference's single-process ``ppo_update`` fixture, and the all-reduced flat gradient must equal the
gradient one GPU computes on the concatenated batch."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_two_rank_ppo_update_matches_single_gpu(cuda_device):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 visible GPUs")
    port = 29500 + os.getpid() % 2000
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", str(port), os.path.join(ROOT, "scripts", "dist_check.py")], cwd=ROOT, capture_output=True,
                       text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count("-> OK") == 2, r.stdout[-3000:]
