"""Experiment (not part of the product): units_per_sm sweep + per-CTA timing of the decode kernel.

    python tools/exp_balance.py [c2|c3|c5]
"""
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
layers = min(layers, 4)
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
combos = os.environ.get("COMBOS", "24:2:0:48,16:2:0:48,20:2:0:48,28:2:0:48,24:2:0:32,24:2:0:64,24:2:0:96,24:2:0:128,"
                                  "16:2:0:96,20:2:0:64,24:1:0:48,16:1:0:64,12:2:0:64")
for combo in combos.split(","):
    ups, gk, st, cl = (int(x) for x in combo.split(":"))
    _lib.check(lib.pa_decode_set_tuning(ups, gk, st, cl), "set_tuning")
    us = timeit()
    res[combo] = us
    print(f"units_per_sm={ups:3d} guided_k={gk} stages={st} claim={cl:3d}  {us:8.2f} us  {kvb / us / 1e3:8.1f} GB/s",
          flush=True)
lib.pa_decode_set_tuning(16, 3, 0, 48)

# per-CTA timestamps
nw = lib.pa_decode_num_workers()
dbg = torch.zeros(nw * 6, dtype=torch.int64, device=dev)
run(20)
torch.cuda.synchronize()
lib.pa_decode_debug_times(dbg.data_ptr())
recs = []
for i in range(4):
    run(1)
    torch.cuda.synchronize()
    recs.append(dbg.cpu().numpy().reshape(nw, 6).copy())
lib.pa_decode_debug_times(None)
r = recs[-1]
t0 = r[:, 1].min()
end = (r[:, 2] - t0) / 1e3
fend = (r[:, 4] - t0) / 1e3
first = (r[:, 5] - t0) / 1e3
start = (r[:, 1] - t0) / 1e3
print(f"per-CTA us: start max {start.max():.1f}; first block med {np.median(first):.1f} max {first.max():.1f}; "
      f"consumers end min {end.min():.1f} med {np.median(end):.1f} max {end.max():.1f}; "
      f"finisher end med {np.median(fend):.1f} max {fend.max():.1f}; "
      f"units per CTA min {r[:, 3].min()} max {r[:, 3].max()}", flush=True)
os.makedirs("gpurun_out", exist_ok=True)
np.savez(f"gpurun_out/balance_{wl_name}.npz", recs=np.stack(recs))
json.dump(res, open(f"gpurun_out/balance_{wl_name}.json", "w"))
