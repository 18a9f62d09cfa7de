"""Multi-GPU parity: sharded filter == one-GPU filter (tests/sharded_worker.py),
launched with torchrun over NCCL.  Needs at least two GPUs; on a one-GPU box the
test is skipped (two ranks must never share a device)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("n,p2p,model,nccl", [(100003, 1, "lf", 1), (4096, 1, "lf", 1), (100003, 0, "lf", 1),
                                              (50001, 1, "lfp", 1), (50001, 1, "lfp_skip", 1),
                                              (20001, 1, "beam", 1), (100003, 1, "lf", 0), (50001, 0, "lfp", 0),
                                              # recovery injection + KLD cut (host callback poses must agree across ranks)
                                              (30001, 1, "lf_recover", 1), (30001, 0, "lf_recover", 0)])
def test_sharded_equals_single_gpu(n, p2p, model, nccl):
    import torch

    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    world = min(torch.cuda.device_count(), 8)
    env = dict(os.environ, SHARD_N=str(n), SHARD_P2P=str(p2p), SHARD_MODEL=model, SHARD_NCCL=str(nccl))
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
                          "--nproc-per-node", str(world), "--master-addr", "127.0.0.1",
                          "--master-port", "29531", os.path.join(ROOT, "tests", "sharded_worker.py")],
                         capture_output=True, text=True, env=env, timeout=900)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    assert "SHARDED-OK" in out.stdout
