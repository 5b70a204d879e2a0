"""Multi-rank parity of KV-head-sharded decode with the fused NVLink output gather.

Spawns ``world = min(8, torch.cuda.device_count())`` ranks (one per GPU, torchrun, NCCL) running
tests/multi_gpu_worker.py - see its docstring for what every rank checks (bit-exact transport for the LL and
counter protocols and the NCCL arm over 50 consecutive steps, ragged lengths incl. seq_len == 0, the C3 per-rank
shape, CUDA-graph replay, dead-peer timeout).  World 2 runs as well when more than 2 GPUs are present (the
odd / even buffer-set alternation and the peer indexing differ from world 8).  The rank logs are kept under
gpurun_out/ (copied to profiles/ by the builder).
"""
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

pytestmark = pytest.mark.gpu


def _worlds():
    n = torch.cuda.device_count() if torch.cuda.is_available() else 0
    return [w for w in (2, 4, 8) if w <= n] or [0]


@pytest.mark.parametrize("world", _worlds())
def test_head_sharded_gather_parity_across_ranks(world):
    if world < 2:
        pytest.skip("needs at least 2 GPUs on the box (the single-GPU kernel side is tests/test_peer_gather.py)")
    log = os.path.join(ROOT, "gpurun_out", f"multi_gpu_parity_w{world}.log")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(29610 + world),
           os.path.join(ROOT, "tests", "multi_gpu_worker.py"), "--steps", "50", "--out", log]
    env = dict(os.environ, NCCL_DEBUG_FILE="/dev/stderr")
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, env=env, cwd=ROOT)
    tail = (r.stdout[-6000:] + "\n--- stderr ---\n" + r.stderr[-3000:])
    assert r.returncode == 0, tail
    assert f"multi_gpu_worker world {world}: ALL OK" in r.stdout, tail


def test_world1_worker_is_bit_equal_to_the_full_head_decode():
    """The same worker with ONE rank: the fused gather path must reproduce the plain decode bit for bit."""
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=1", "--master-addr", "127.0.0.1",
           "--master-port", "29609", os.path.join(ROOT, "tests", "multi_gpu_worker.py"), "--steps", "10"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT,
                       env=dict(os.environ, CUDA_VISIBLE_DEVICES=os.environ.get("CUDA_VISIBLE_DEVICES", "0").split(",")[0]))
    assert r.returncode == 0 and "world 1: ALL OK" in r.stdout, r.stdout[-4000:] + r.stderr[-2000:]
