"""Secondary metrics of the path ("KV write and block ops"): the slot write
(reshape-and-cache), block copy and block swap, each against the roofline that bounds it.

    python tools/bench_cache_ops.py [--iters 50]

One JSON line per op.  Shapes are the C2 layer: Hkv 8, D 128, block_size 16, bf16, pool
[2, NB, Hkv, BS, D] ('hbsd', the reference pool of 01_paged_attention/paged_attention.py:26-30) and
the engine layout (k_cache, v_cache)[NB, BS, H, D] ('bshd', 05_mini_vllm/mini_vllm.py:544-552).

  * reshape_and_cache  T = 64 (the decode step: latency-bound, reported in microseconds) and
    T = 8192 (a prefill chunk): algorithmic bytes = 2*T*Hkv*D*2 read + the same written + 8*T slots.
  * copy_blocks        N = 1024 (src, dst) pairs: 2 * 64 KiB per pair (read + write).
  * swap_out / swap_in N = 1024 blocks (64 MiB) between the device pool and a pinned host pool;
    the roofline is the host link: a plain pinned cudaMemcpyAsync of the same bytes, measured here
    first, is the peak the fraction is quoted against (not HBM).

Timing: CUDA events on the launching stream, median of --iters after warm-up; every op works on a
pool larger than L2 (NB = 16384 blocks = 1 GiB per layer) with fresh random slots / pairs.
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vllm_from_scratch_b200 import paged_attention as P            # noqa: E402

HKV, D, BS = 8, 128, 16
NB = 16384


def hbm_peak():
    try:
        m = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(m["hbm_gbs"]), "measured copy bandwidth (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:                                               # noqa: BLE001
        return 6553.0, "fallback (B200_PROFILING.md)"


def timed(fn, iters, warmup=5, stream=None, setup=None):
    """Median / min milliseconds of fn() between two events on `stream` (default: current)."""
    st = stream or torch.cuda.current_stream()
    for _ in range(warmup):
        if setup:
            setup()
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if setup:
            setup()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        fn()
        e1.record(st)
        e1.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2], ts[0]


def timed_graph(fns, iters, warmup=3):
    """Device time per call with the host launch path taken out: the calls in `fns` (each on its
    own inputs, together larger than L2) are captured into ONE CUDA graph; returns the median /
    min milliseconds of a replay divided by len(fns)."""
    for f in fns:
        f()
    torch.cuda.synchronize()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr):
        for f in fns:
            f()
    for _ in range(warmup):
        gr.replay()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        gr.replay()
        e1.record()
        e1.synchronize()
        ts.append(e0.elapsed_time(e1) / len(fns))
    ts.sort()
    return ts[len(ts) // 2], ts[0]


def line(op, shape, nbytes, ms, ms_min, peak, unit_peak, src, bound, **extra):
    gbs = nbytes / (ms * 1e-3) / 1e9
    d = {"op": op, "shape": shape, "algorithmic_bytes": nbytes, "us": ms * 1e3, "us_min": ms_min * 1e3,
         "roofline": {"bound": bound, "achieved": gbs, "peak": peak, "unit": unit_peak, "frac": gbs / peak,
                      "peak_source": src}}
    d.update(extra)
    print(json.dumps(d), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=50)
    a = ap.parse_args()
    dev = "cuda"
    hbm, hbm_src = hbm_peak()
    g = torch.Generator(device=dev).manual_seed(0)
    cg = torch.Generator().manual_seed(0)
    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2

    def flush():
        flush_buf.zero_()

    pool = torch.randn(2, NB, HKV, BS, D, generator=g, device=dev, dtype=torch.float32).to(torch.bfloat16)
    k_cache = torch.zeros(NB, BS, HKV, D, device=dev, dtype=torch.bfloat16)
    v_cache = torch.zeros_like(k_cache)
    block_bytes = 2 * HKV * BS * D * 2                                    # K + V of one block, one layer

    # ---- slot write ----------------------------------------------------------------------
    # `us` = device time per call inside a CUDA graph of R calls on R different inputs (host launch
    # path excluded, inputs + written slots of a replay larger than L2); `api_us` = one eager call
    # through the Python API, events around it (includes the host-side launch path).
    for t in (64, 8192):
        reps = 8
        ks = [torch.randn(t, HKV, D, generator=g, device=dev, dtype=torch.float32).to(torch.bfloat16) for _ in range(reps)]
        vs = [torch.randn(t, HKV, D, generator=g, device=dev, dtype=torch.float32).to(torch.bfloat16) for _ in range(reps)]
        scat, runs = [], []
        for _ in range(reps):
            scat.append(torch.randperm(NB * BS, generator=cg)[:t].to(dev))       # scattered single-token slots
            # a prefill chunk writes consecutive slots inside each block: 16-token runs in random blocks
            blocks = torch.randperm(NB, generator=cg)[:(t + BS - 1) // BS]
            runs.append((blocks[:, None] * BS + torch.arange(BS)[None, :]).reshape(-1)[:t].to(dev))
        nbytes = 2 * (2 * t * HKV * D * 2) + 8 * t
        for layout, cache in (("hbsd", pool), ("bshd", (k_cache, v_cache))):
            for kind, ss in (("scattered", scat), ("block-runs", runs)):
                fns = [(lambda i=i: P.reshape_and_cache(ks[i], vs[i], cache, ss[i])) for i in range(reps)]
                ms, mn = timed_graph(fns, a.iters)
                api, _ = timed(fns[0], a.iters, setup=flush)
                line("reshape_and_cache", f"T={t} layout={layout} slots={kind}", nbytes, ms, mn, hbm, "GB/s", hbm_src,
                     "hbm" if t > 1024 else "latency", api_us=api * 1e3)

    # ---- block copy ------------------------------------------------------------------------
    n = 1024
    reps = 4
    perm = torch.randperm(NB, generator=cg)
    pairs = [torch.stack([perm[2 * i * n:(2 * i + 1) * n], perm[(2 * i + 1) * n:(2 * i + 2) * n]], dim=1).to(torch.int64).to(dev)
             for i in range(reps)]
    for label, cache in ((f"N={n} blocks layout=hbsd (2 planes of 32 KiB per block)", pool),
                         (f"N={n} blocks layout=bshd (k_cache, v_cache; 32 KiB per block each)", (k_cache, v_cache))):
        fns = [(lambda i=i: P.copy_blocks(cache, pairs[i])) for i in range(reps)]
        ms, mn = timed_graph(fns, a.iters)
        api, _ = timed(fns[0], a.iters, setup=flush)
        line("copy_blocks", label, 2 * n * block_bytes, ms, mn, hbm, "GB/s", hbm_src, "hbm", api_us=api * 1e3)

    # ---- swap ------------------------------------------------------------------------------
    side = torch.cuda.Stream()
    nbytes = n * block_bytes
    host_flat = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    dev_flat = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    peaks = {}
    for name, dst, src in (("d2h", host_flat, dev_flat), ("h2d", dev_flat, host_flat)):
        with torch.cuda.stream(side):
            ms, mn = timed(lambda: dst.copy_(src, non_blocking=True), max(a.iters // 2, 5), stream=side)
        peaks[name] = nbytes / (ms * 1e-3) / 1e9
        line(f"pinned_memcpy_{name}", f"{nbytes >> 20} MiB contiguous cudaMemcpyAsync (the link peak)", nbytes, ms, mn,
             peaks[name], "GB/s", "this measurement", "host-link")
    host = P.HostKVPool(pool, n)
    scattered = perm[:n].tolist()
    runs = list(range(4096, 4096 + n))
    for kind, ids in (("scattered blocks", scattered), ("one run of consecutive blocks", runs)):
        out_pairs = [(b, i) for i, b in enumerate(ids)]
        in_pairs = [(i, b) for i, b in enumerate(ids)]
        ms, mn = timed(lambda: P.swap_out(pool, host.kv_cache, out_pairs, stream=side), max(a.iters // 2, 5), stream=side)
        line("swap_out", f"N={n} blocks, {kind}", nbytes, ms, mn, peaks["d2h"], "GB/s",
             "pinned cudaMemcpyAsync D2H of the same bytes, measured in this run", "host-link")
        ms, mn = timed(lambda: P.swap_in(host.kv_cache, pool, in_pairs, stream=side), max(a.iters // 2, 5), stream=side)
        line("swap_in", f"N={n} blocks, {kind}", nbytes, ms, mn, peaks["h2d"], "GB/s",
             "pinned cudaMemcpyAsync H2D of the same bytes, measured in this run", "host-link")


if __name__ == "__main__":
    main()
