"""Experiment (not part of the product): where the in-kernel cost of the fused gather goes.  Under
torchrun (2+ ranks): times pa_decode_gather on the head-sharded C3 shape with the remote stores and
the remote counters redirected to LOCAL scratch memory one at a time."""
import copy
import ctypes
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from vllm_from_scratch_b200 import _lib, paged_attention as P, sharding as S  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
b, ctx, d, bs, L = 8, 32768, 128, 16, 4
hkv, hq = 8 // world, 32 // world
pool, bts_h, lens_h, q_h, *_ = bench.make_workload(torch, dev, (b, ctx, hq, hkv, d, bs, L, ""), L, seed=1)
bts, lens, q = bts_h.to(dev), lens_h.to(dev), q_h.to(dev)
ws = P.new_decode_workspace(b, hq, hkv, d, dev)
caches = [pool.layer(i) for i in range(L)]
pg = S.PeerGather(b, 32, 8, d, dev, world=world, rank=rank)
scratch_out = torch.zeros(b * hq * d, dtype=torch.bfloat16, device=dev)
scratch_sig = torch.zeros(64, dtype=torch.int32, device=dev)
out = pg.sets[0][rank]


def variant(stores_remote, sig_remote, nremote=None):
    src = pg._peers[0]
    v = _lib.PaPeerGather()
    v.num_remote = src.num_remote if nremote is None else nremote
    for j in range(7):
        v.remote_out[j] = src.remote_out[j] if stores_remote else scratch_out.data_ptr()
        v.remote_sig[j] = src.remote_sig[j] if sig_remote else scratch_sig.data_ptr() + 4 * j
    v.local_sig = scratch_sig.data_ptr() + 32 * 4
    return v


def timeit(peers, n=300):
    def run(k):
        for i in range(k):
            P.paged_attention_decode_batched(q, caches[i % L], bts, lens, out=out, workspace=ws, peers=peers)
    run(30)
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    run(n)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


res = {}
for name, peers in [("plain pa_decode", None), ("gather, 0 remote (local counter only)", variant(False, False, 0)),
                    ("stores local scratch, counters local scratch", variant(False, False)),
                    ("stores REMOTE, counters local", variant(True, False)),
                    ("stores local, counters REMOTE", variant(False, True)),
                    ("stores REMOTE, counters REMOTE", variant(True, True))]:
    res[name] = timeit(peers)
    dist.barrier()
if rank == 0:
    for k, v in res.items():
        print(f"{k:50s} {v:7.1f} us", flush=True)
torch.cuda.synchronize()
dist.barrier()
pg.close()
dist.destroy_process_group()
