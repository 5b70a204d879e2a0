"""Experiment (not part of the product): split-K plan sweep + per-SM timing of the decode kernel."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from vllm_from_scratch_b200 import _lib, paged_attention as P  # noqa: E402

wl_name = sys.argv[1] if len(sys.argv) > 1 else "c2"
wl = bench.WORKLOADS[wl_name]
b, ctx, hq, hkv, d, bs, layers, _ = wl
layers = 4
dev = torch.device("cuda", 0)
lib = _lib.load()
pool, bts_h, lens_h, q_h, *_ = bench.make_workload(torch, dev, wl, layers, seed=1)
bts, lens, q = bts_h.to(dev), lens_h.to(dev), q_h.to(dev)
out = torch.empty_like(q)
ws = P.new_decode_workspace(b, hq, hkv, d, dev)
caches = [pool.layer(i) for i in range(layers)]
kvb, iob = bench.algorithmic_bytes(b, ctx, hq, hkv, d, bs)


def run(n):
    for i in range(n):
        P.paged_attention_decode_batched(q, caches[i % layers], bts, lens, out=out, workspace=ws)


def timeit(n=300):
    run(30)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    run(n)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


run(500)
res = {}
settings = [(0, 4), (4, 4), (8, 4), (8, 8), (16, 4), (16, 2), (16, 8), (32, 4), (32, 2), (6, 16), (8, 16)]
for pd, cm in settings:
    lib.pa_decode_set_tuning(pd, cm)
    us = timeit()
    res[f"{pd},{cm}"] = us
    print(f"pool_div={pd:3d} chunk_min={cm:3d}  {us:8.2f} us  {kvb / us / 1e3:8.1f} GB/s", flush=True)

# per-worker timestamps
nw = lib.pa_decode_num_workers()
dbg = torch.zeros(nw * 6, dtype=torch.int64, device=dev)
allt = {}
for pd, cm in [(0, 4), (8, 4)]:
    lib.pa_decode_set_tuning(pd, cm)
    run(20)
    torch.cuda.synchronize()
    lib.pa_decode_debug_times(dbg.data_ptr())
    recs = []
    for i in range(4):
        run(1)
        torch.cuda.synchronize()
        recs.append(dbg.cpu().numpy().reshape(nw, 6).copy())
    lib.pa_decode_debug_times(None)
    allt[f"{pd}_{cm}"] = np.stack(recs)
np.savez(f"gpurun_out/balance_{wl_name}.npz", **allt)
json.dump(res, open(f"gpurun_out/balance_{wl_name}.json", "w"))
lib.pa_decode_set_tuning(4, 4)
