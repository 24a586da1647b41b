"""Multi-GPU parity: the point-sharded reduced camera system and LM equal the unsharded ones.

Needs >= 2 GPUs on the box; skipped otherwise (the host-side shard logic is covered on CPU by test_dist_gloo.py).
"""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _gpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.gpu
@pytest.mark.parametrize("xchg", ["0", "1", "p2p"])
@pytest.mark.parametrize("shape", [(40, 4000, 16000), (300, 30000, 120000)])
def test_sharded_equals_unsharded(xchg, shape):
    """xchg: 0 = all-reduce of the packed buffer, 1 = sparse exchange through ncclAllGather, p2p = sparse exchange through
    peer memory (CUDA-IPC mapped receive slots, k_p2p_push / k_p2p_wait) -- all three must equal the unsharded system."""
    n = _gpus()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    world = 4 if n >= 4 else 2
    env = dict(os.environ, B200BA_XCHG="1" if xchg == "p2p" else xchg, B200BA_P2P="1" if xchg == "p2p" else "0")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
           "127.0.0.1", "--master-port", "29533", os.path.join(ROOT, "tests", "_multi_worker.py")] + [str(v) for v in shape]
    p = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
    assert p.stdout.count(" OK") == world, p.stdout
