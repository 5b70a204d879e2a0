"""Experiment (not part of the product): the two generic decode kernels (cp.async streaming / plain) on the C2 batch
with an fp32 and a bf16 cache, graph-timed.   python tools/exp_generic.py [f32|bf16] [iters]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from vllm_from_scratch_b200 import _lib, paged_attention as P  # noqa: E402

dt = torch.float32 if (len(sys.argv) < 2 or sys.argv[1] == "f32") else torch.bfloat16
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 40
dev = torch.device("cuda", 0)
lib = _lib.load()
b, ctx, hq, hkv, d, bs, _, _ = bench.WORKLOADS["c2"]
layers = 2
nblk = ctx // bs
nb = b * nblk + 64
pool = P.KVCachePool(nb, bs, hkv, d, layers, dtype=dt, device=dev)
pool.kv_cache.normal_()
gc = torch.Generator().manual_seed(5)
bts = torch.randperm(nb, generator=gc)[: b * nblk].to(torch.int32).view(b, nblk).to(dev)
lens = torch.full((b,), ctx, dtype=torch.int32, device=dev)
q = torch.randn(b, hq, d, generator=gc).to(dt).to(dev)
out = torch.empty_like(q)
ws = P.new_decode_workspace(b, hq, hkv, d, dev)
caches = [pool.layer(i) for i in range(layers)]
kvb, iob = bench.algorithmic_bytes(b, ctx, hq, hkv, d, bs, esize=pool.kv_cache.element_size())
for path, name in ((1, "stream"), (2, "plain")):
    lib.pa_decode_force_path(path)
    graphs = bench.capture_steps(torch, lambda i: P.paged_attention_decode_batched(q, caches[i], bts, lens, out=out, workspace=ws), layers)
    for i in range(4):
        graphs[i % layers].replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        graphs[i % layers].replay()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / iters * 1e3
    print(f"{name:7s} {str(dt):15s} {us:8.1f} us  {(kvb + iob) / us / 1e3:7.0f} GB/s", flush=True)
lib.pa_decode_force_path(0)
