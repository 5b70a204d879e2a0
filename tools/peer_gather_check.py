"""Multi-rank parity of the fused head-sharded output gather (run under torchrun, one rank per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tools/peer_gather_check.py

Every rank builds the SAME full batch (same seed), keeps its KV-head slice, runs PeerGather.decode for
several steps (rank-dependent delays in between so the ranks drift apart) and checks the gathered
[B, Hq, D] (max-abs 1e-2; the split-K merge order depends on the local head count) against the full-head decode it computes locally.  Prints one line per rank."""
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vllm_from_scratch_b200 import paged_attention as P  # noqa: E402
from vllm_from_scratch_b200 import sharding as S  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    hq, hkv, d, bs = 32, 8, 128, 16
    lens = [0, 1, 17, 5000, 333, 40000, 16, 2048]
    g = torch.Generator().manual_seed(7)
    sl = torch.tensor(lens, dtype=torch.int32)
    nblk = (sl + bs - 1) // bs
    nb = int(nblk.sum()) + 8
    kv = torch.randn(2, nb, hkv, bs, d, generator=g).to(torch.bfloat16).to(dev)
    perm = torch.randperm(nb, generator=g).to(torch.int32)
    bts = torch.zeros(len(lens), int(nblk.max()), dtype=torch.int32)
    at = 0
    for i, n in enumerate(nblk.tolist()):
        bts[i, :n] = perm[at:at + n]
        at += n
    bts, sl = bts.to(dev), sl.to(dev)
    kvh, qh = S.shard_kv_heads(hkv, world, rank), S.shard_q_heads(hq, hkv, world, rank)
    kv_local = kv[:, :, kvh.start:kvh.stop].contiguous()
    ok = True
    for protocol in ("ll", "counter"):
        pg = S.PeerGather(len(lens), hq, hkv, d, dev, world=world, rank=rank, timeout_us=5_000_000, protocol=protocol)
        for step in range(6):
            q = torch.randn(len(lens), hq, d, generator=g).to(torch.bfloat16).to(dev)     # same on every rank
            want = P.paged_attention_decode_batched(q, kv, bts, sl)
            time.sleep(0.002 * ((rank + step) % world))                                   # let the ranks drift
            got = pg.full(pg.decode(q[:, qh.start:qh.stop].contiguous(), kv_local, bts, sl))
            pg.check()
            # split-K unit boundaries depend on the number of local heads, so merge order (not values) differs
            err = (got.float() - want.float()).abs().max().item()
            ok &= err <= 1e-2
            if err > 1e-2:
                print(f"rank {rank} {protocol} step {step}: max diff {err}", flush=True)
        pg.close()
    print(f"peer_gather_check rank {rank}/{world}: {'OK' if ok else 'MISMATCH'}", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
