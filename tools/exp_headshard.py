"""Experiment (not part of the product): the decode kernel on the PER-RANK shapes of KV-head-sharded
C3 (8 seqs x 32k, Hkv / P heads), on ONE GPU: graph-timed kernel (no host launch path), tuning sweep,
per-CTA timestamps, and the cost of the fused-gather signalling (world-1 PeerGather).

    python tools/exp_headshard.py
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from vllm_from_scratch_b200 import _lib, paged_attention as P, sharding as S  # noqa: E402

dev = torch.device("cuda", 0)
lib = _lib.load()
LAYERS = 4


def graph_time(fn, reps=LAYERS, iters=60):
    for i in range(reps):
        fn(i)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(reps):
            fn(i)
    for _ in range(5):
        g.replay()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        e1.synchronize()
        ts.append(e0.elapsed_time(e1) / reps * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


for world in [int(x) for x in os.environ.get("WORLDS", "1,2,4,8").split(",")]:
    b, ctx, d, bs = 8, 32768, 128, 16
    hkv, hq = 8 // world, 32 // world
    wl = (b, ctx, hq, hkv, d, bs, LAYERS, "")
    pool, bts_h, lens_h, q_h, *_ = bench.make_workload(torch, dev, wl, LAYERS, seed=1)
    bts, lens, q = bts_h.to(dev), lens_h.to(dev), q_h.to(dev)
    out = torch.empty_like(q)
    ws = P.new_decode_workspace(b, hq, hkv, d, dev)
    caches = [pool.layer(i) for i in range(LAYERS)]
    kvb, _ = bench.algorithmic_bytes(b, ctx, hq, hkv, d, bs)
    ideal = kvb / 6.55e12 * 1e6
    print(f"== P={world}: {hkv} KV heads / rank, {kvb >> 20} MiB, ideal {ideal:.1f} us at 6.55 TB/s", flush=True)
    for combo in os.environ.get("COMBOS", "16:3:0:48,24:2:0:48,32:2:0:48,32:4:0:48,48:4:0:48,64:4:0:32,64:8:0:32").split(","):
        ups, gk, st, cl = (int(x) for x in combo.split(":"))
        if lib.pa_decode_set_tuning(ups, gk, st, cl) != 0:
            continue
        us = graph_time(lambda i: P.paged_attention_decode_batched(q, caches[i], bts, lens, out=out, workspace=ws))
        print(f"   units_per_sm={ups:3d} guided_k={gk} claim={cl:3d}: {us:7.1f} us  {kvb / us / 1e3:7.0f} GB/s", flush=True)
    lib.pa_decode_set_tuning(16, 3, 0, 32)
    if os.environ.get("NO_PEER"):
        continue
    pg = S.PeerGather(b, hq, hkv, d, dev, world=1, rank=0)
    us = graph_time(lambda i: pg.decode(q, caches[i], bts, lens, workspace=ws, parity=i % 2))
    print(f"   fused gather (world 1: local counters + wait kernel): {us:7.1f} us", flush=True)
    pg.check()
    pg.close()
    # per-CTA timestamps of one launch
    nw = lib.pa_decode_num_workers()
    dbg = torch.zeros(nw * 16, dtype=torch.int64, device=dev)
    lib.pa_decode_debug_times(dbg.data_ptr())
    for i in range(3):
        P.paged_attention_decode_batched(q, caches[i], bts, lens, out=out, workspace=ws)
        torch.cuda.synchronize()
    lib.pa_decode_debug_times(None)
    r = dbg.cpu().numpy().reshape(nw, 16)
    t0 = r[:, 1].min()
    end, fend, first = (r[:, 2] - t0) / 1e3, (r[:, 4] - t0) / 1e3, (r[:, 5] - t0) / 1e3
    print(f"   per-CTA us: first block med {np.median(first):.1f} max {first.max():.1f}; consumers end min {end.min():.1f} "
          f"med {np.median(end):.1f} max {end.max():.1f}; finisher end med {np.median(fend):.1f} max {fend.max():.1f}; "
          f"units/CTA {r[:, 3].min()}..{r[:, 3].max()}", flush=True)
    del pool, caches
    torch.cuda.empty_cache()
