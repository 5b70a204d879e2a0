"""Experiment (not part of the product): split-K tuning sweep + per-CTA timeline of the decode kernel.

    python tools/exp_balance.py [c2|c3|c5] [--combos ups:gk:stages:claim,...] [--out gpurun_out/timeline_c2.json]

Prints, per tuning combo, the device time per launch (CUDA graph replays over the layer caches, so the host
launch path is out of the picture) and the per-CTA timeline of one launch taken with
pa_decode_debug_times (include/pa_b200_debug.h): plan built, first unit published, first TMA load issued,
first KV block landed, tickets exhausted, consumers' end, finisher's end - all relative to the earliest CTA start.
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from vllm_from_scratch_b200 import _lib, paged_attention as P  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("workload", nargs="?", default="c2")
ap.add_argument("--combos", default="16:3:0:48")
ap.add_argument("--out", default="")
ap.add_argument("--iters", type=int, default=400)
ap.add_argument("--pdl", type=int, default=1)
ap.add_argument("--one-graph", type=int, default=1)
args = ap.parse_args()

wl = bench.WORKLOADS[args.workload]
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
W = 16


def launch(i):
    P.paged_attention_decode_batched(q, caches[i % layers], bts, lens, out=out, workspace=ws)


def timeit(n):
    # the tuning knobs are baked in at capture time
    graphs = bench.capture_steps(torch, launch, layers, one_graph=bool(args.one_graph))
    per = layers if args.one_graph else 1
    for i in range(40 // per):
        graphs[i % len(graphs)].replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n // per):
        graphs[i % len(graphs)].replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (n // per * per) * 1e3


def timeline():
    nw = lib.pa_decode_num_workers()
    dbg = torch.zeros(nw * W, dtype=torch.int64, device=dev)
    for i in range(8):
        launch(i)
    torch.cuda.synchronize()
    lib.pa_decode_debug_times(dbg.data_ptr())
    recs = []
    for i in range(5):
        launch(i)
        torch.cuda.synchronize()
        recs.append(dbg.cpu().numpy().reshape(nw, W).copy())
    lib.pa_decode_debug_times(None)
    r = recs[-1]
    t0 = r[:, 1].min()
    rel = lambda col: (r[:, col] - t0) / 1e3                        # noqa: E731
    pct = lambda x: [round(float(v), 2) for v in np.percentile(x, [0, 50, 100])]     # noqa: E731
    return {"cta_start": pct(rel(1)), "plan_built": pct(rel(6)), "first_unit_published": pct(rel(7)),
            "first_tma_issued": pct(rel(8)), "first_block_landed": pct(rel(5)), "tickets_exhausted": pct(rel(9)),
            "consumers_end": pct(rel(2)), "finisher_end": pct(rel(4)),
            # cycles -> us at 1.965 GHz: time consumer warp 0 waited for unit descriptors / for KV data / for the
            # hand-off area to the finishers
            "cons_wait_desc_us": pct(r[:, 13] / 1965.0), "cons_wait_data_us": pct(r[:, 14] / 1965.0),
            "cons_wait_handoff_us": pct(r[:, 15] / 1965.0),
            "units_per_cta": [int(r[:, 3].min()), int(np.median(r[:, 3])), int(r[:, 3].max())],
            # the chain behind the slowest finishers: last unit merged, last fold job taken, its last batch seen, done
            "slowest_finishers": [{"cta": int(i), "consumers_end": round(float(rel(2)[i]), 2),
                                   "last_merge": round(float(rel(10)[i]), 2), "last_job_taken": round(float(rel(11)[i]), 2),
                                   "last_batch_seen": round(float(rel(12)[i]), 2), "finisher_end": round(float(rel(4)[i]), 2)}
                                  for i in np.argsort(-rel(4))[:4]]}, np.stack(recs)


lib.pa_decode_set_pdl(args.pdl)
res = {}
for combo in args.combos.split(","):
    ups, gk, st, cl = (int(x) for x in combo.split(":"))
    _lib.check(lib.pa_decode_set_tuning(ups, gk, st, cl), "set_tuning")
    us = timeit(args.iters)
    tl, recs = timeline()
    res[combo] = {"us_per_launch": us, "kv_gbs": kvb / us / 1e3, "timeline_us_min_med_max": tl}
    print(f"units_per_sm={ups:3d} guided_k={gk} stages={st} claim={cl:3d}  {us:8.2f} us  {kvb / us / 1e3:8.1f} GB/s", flush=True)
    for k, v in tl.items():
        print(f"      {k:22s} {v}", flush=True)
lib.pa_decode_set_tuning(16, 3, 0, 32)
if args.out:
    os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
    json.dump({"workload": args.workload, "results": res}, open(args.out, "w"), indent=1)
