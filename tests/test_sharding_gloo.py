"""Multi-GPU partitioning on CPU: world_size 2 over gloo.

The kernels are rank-local; what needs a multi-process test is the partition arithmetic and the
one collective (the KV-head-sharded output gather).  Each rank computes its shard's attention with
the CPU oracle standing in for the kernel and the gathered result must equal the unsharded one."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from vllm_from_scratch_b200 import sharding as S  # noqa: E402


def test_shard_arithmetic():
    for b, w in [(64, 8), (512, 8), (7, 4), (3, 8), (0, 2)]:
        got = [i for r in range(w) for i in S.shard_sequences(b, w, r)]
        assert got == list(range(b))
        sizes = [len(S.shard_sequences(b, w, r)) for r in range(w)]
        assert max(sizes) - min(sizes) <= 1
    assert [list(S.shard_kv_heads(8, 4, r)) for r in range(4)] == [[0, 1], [2, 3], [4, 5], [6, 7]]
    assert list(S.shard_q_heads(32, 8, 4, 1)) == list(range(8, 16))
    with pytest.raises(ValueError):
        S.shard_kv_heads(8, 3, 0)
    with pytest.raises(ValueError):
        S.shard_sequences(4, 2, 2)


def _worker(rank, world, port, mode, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import paged_attention_oracle as O
        g = torch.Generator().manual_seed(7)              # same data on every rank
        b, hq, hkv, d, bs, nb = 5, 8, 4, 16, 4, 40
        kv = torch.randn(2, nb, hkv, bs, d, generator=g)
        q = torch.randn(b, hq, d, generator=g)
        lens = torch.tensor([13, 1, 24, 7, 16], dtype=torch.int32)
        perm = torch.randperm(nb, generator=g).to(torch.int32)
        bts = perm[:b * 6].view(b, 6)
        full = O.decode_batched(q, kv, bts, lens, bs)
        if mode == "head":
            kvh = S.shard_kv_heads(hkv, world, rank)
            qh = S.shard_q_heads(hq, hkv, world, rank)
            gather = S.HeadGather(b, hq, d, torch.float32, "cpu")
            local = O.decode_batched(q[:, qh.start:qh.stop], kv[:, :, kvh.start:kvh.stop].contiguous(), bts, lens, bs)
            gather.local_out().copy_(local)               # the kernel writes here directly on the GPU
            gather.all_gather()
            got = gather.full()
            assert torch.equal(gather.full_view()[:, rank], local)
        else:
            mine = S.shard_sequences(b, world, rank)
            idx = list(mine)
            local = O.decode_batched(q[idx], kv, bts[idx], lens[idx], bs) if idx else q[:0]
            got = S.gather_sequence_outputs(local, b, world, rank)
        ok = torch.allclose(got, full, atol=1e-6)
        flag = torch.tensor([1 if ok else 0])
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if rank == 0:
            ret.put(int(flag.item()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("mode", ["head", "seq"])
def test_sharded_decode_world2_gloo(mode):
    ctx = mp.get_context("spawn")
    ret = ctx.Queue()
    port = 29600 + (os.getpid() % 200) + (0 if mode == "head" else 211)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, mode, ret)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert ret.get(timeout=5) == 1
