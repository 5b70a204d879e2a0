#!/usr/bin/env python
"""bench.py - paged-attention decode tokens/s and KV-read GB/s on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2|c3|c5]

A "step" is one pass of the hot path over one decode batch: ONE attention-layer call
(one launch: split-K decode kernel that merges its own partials) for every sequence of the batch.  At N = 1
the workload is BASELINE.json configs[1] ("C2": 64 seqs x 2k context, 32 Q / 8 KV heads,
head_dim 128, bf16, block_size 16).  For N > 1 the batch is sharded by sequence: every rank
owns its own block pool and block tables and runs the same per-GPU batch (weak scaling, no
data-path collective); value = sequences all ranks advanced / max-over-ranks device time.

L2: the timed loop rotates over `--layers` per-layer caches of the pool (C2: 8 x 512 MiB =
4 GiB >> 126 MB L2), i.e. inputs larger than L2; nothing is re-read from cache.

One JSON line on stdout (rank 0).  `--impl reference` times the CPU oracle port of the
reference's own algorithm (oracle/paged_attention_oracle.py::decode_loop) on the host cores.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "paged_attn_decode_tokens_per_s"
UNIT = "tokens/s"
NOMINAL_HBM_GBS = 8000.0          # the metric is quoted against 8 TB/s (BASELINE.json)
FALLBACK_HBM_GBS = 6650.0         # /opt/skills/guides/B200_PROFILING.md fallback

WORKLOADS = {
    # name: (seqs per GPU, context, Hq, Hkv, D, BS, default layers, description)
    "c2": (64, 2048, 32, 8, 128, 16, 8,
           "C2 Llama-3-8B-shaped decode: 64 seqs x 2048 ctx, 32 Q / 8 KV heads, head_dim 128, bf16, block_size 16"),
    "c3": (8, 32768, 32, 8, 128, 16, 4,
           "C3 long-context decode: 8 seqs x 32768 ctx, GQA 32/8, head_dim 128, bf16, block_size 16"),
    "c5": (64, 8192, 32, 8, 128, 16, 2,
           "C5 shard: 64 seqs x 8192 ctx per GPU (512 seqs over 8 GPUs), GQA 32/8, head_dim 128, bf16"),
    # shape scan (experiments only; same KV byte count as C2 unless noted)
    "x_mha": (16, 2048, 32, 32, 128, 16, 4, "MHA: 16 seqs x 2048 ctx, 32 Q / 32 KV heads"),
    "x_g8": (64, 2048, 64, 8, 128, 16, 4, "GQA 8: 64 seqs x 2048 ctx, 64 Q / 8 KV heads"),
    "x_b1": (1, 131072, 32, 8, 128, 16, 4, "one sequence of 131072 tokens"),
    "x_b512": (512, 256, 32, 8, 128, 16, 4, "512 seqs x 256 ctx"),
    "x_b4": (4, 4096, 32, 8, 128, 16, 4, "4 seqs x 4096 ctx (32 MiB of KV)"),
    "x_b1s": (1, 2048, 32, 8, 128, 16, 4, "1 seq x 2048 ctx (4 MiB of KV)"),
    "x_b8": (8, 1024, 32, 8, 128, 16, 4, "8 seqs x 1024 ctx (16 MiB of KV)"),
    "c2d64": (64, 4096, 32, 8, 64, 16, 8,
              "C2-sized decode at head_dim 64: 64 seqs x 4096 ctx, 32 Q / 8 KV heads, bf16 (experiments only)"),
    "c3r8": (8, 32768, 4, 1, 128, 16, 4,
             "per-rank shape of KV-head-sharded C3 at 8 GPUs: 8 seqs x 32768 ctx, 4 Q / 1 KV head (experiments only)"),
    "c5h": (64, 32768, 32, 8, 128, 16, 2,
            "C5 32k shape: 64 seqs x 32768 ctx (8 GiB of KV per layer), GQA 32/8, head_dim 128, bf16 - the KV-head-"
            "sharded whole-box case (--shard head)"),
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=50)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=None, choices=sorted(WORKLOADS))
    ap.add_argument("--shard", default="seq", choices=["seq", "head"],
                    help="N > 1: 'seq' = every rank its own sequences (weak scaling, no collective); 'head' = "
                         "every rank Hkv / N KV heads of the SAME batch + one NCCL all-gather of the outputs "
                         "(strong scaling; default workload c3)")
    ap.add_argument("--gather", default="peer", choices=["peer", "peer-counter", "nccl"],
                    help="--shard head: 'peer' = output gather fused into the decode kernel over NVLink peer memory "
                         "(default); 'nccl' = decode kernel + one NCCL all-gather")
    ap.add_argument("--layers", type=int, default=0, help="per-layer caches the loop rotates over")
    ap.add_argument("--e2e-steps", type=int, default=0, help="steps of the host-buffer leg (default min(steps, 300))")
    ap.add_argument("--ramp-ms", type=float, default=300.0, help="untimed clock ramp before the warm-up steps")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true",
                    help="skip the extra driver-visible points of the default run (c5_512x8k, generic_f32, head_sharded)")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="budget of the cpu_baseline sample")
    return ap.parse_args()


# ---------------------------------------------------------------------------------------
# clocks sampling during the timed region
# ---------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU every few ms through NVML."""

    REASONS = {
        0x0000000000000004: "sw_power_cap",
        0x0000000000000008: "hw_slowdown",
        0x0000000000000020: "sw_thermal_slowdown",
        0x0000000000000040: "hw_thermal_slowdown",
        0x0000000000000080: "hw_power_brake_slowdown",
    }

    def __init__(self, index: int, period_s: float = 0.002):
        super().__init__(daemon=True)
        self.index = index
        self.period = period_s
        self.samples = []          # (t, sm_mhz, reasons_mask, power_w)
        self.sm_max = None
        self.stop_flag = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(self._physical_index(index))
            self.sm_max = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception as e:      # noqa: BLE001 - clocks are advisory; report the failure in JSON
            self.err = repr(e)

    @staticmethod
    def _physical_index(i: int) -> int:
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            ids = [x for x in vis.split(",") if x.strip() != ""]
            try:
                return int(ids[i])
            except (ValueError, IndexError):
                return i
        return i

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        while not self.stop_flag.is_set():
            try:
                mhz = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                try:
                    mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:   # noqa: BLE001
                    mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                try:
                    pw = nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0
                except Exception:   # noqa: BLE001
                    pw = None
                self.samples.append((time.perf_counter(), mhz, mask, pw))
            except Exception:       # noqa: BLE001
                pass
            time.sleep(self.period)

    def summary(self, t0: float, t1: float) -> dict:
        if not self.ok:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "error": getattr(self, "err", "nvml unavailable")}
        inside = [s for s in self.samples if t0 <= s[0] <= t1]
        where = "timed_region"
        if not inside:              # region shorter than the sampling period: nearest samples
            inside = sorted(self.samples, key=lambda s: abs(s[0] - t1))[:3]
            where = "nearest_to_timed_region"
        mask = 0
        for s in inside:
            mask |= s[2]
        reasons = [name for bit, name in self.REASONS.items() if mask & bit]
        pws = [s[3] for s in inside if s[3] is not None]
        return {"sm_mhz": statistics.median(s[1] for s in inside) if inside else None,
                "sm_max_mhz": self.sm_max, "reasons": reasons, "samples": len(inside), "where": where,
                "power_w_max": max(pws) if pws else None}


# ---------------------------------------------------------------------------------------
# workload
# ---------------------------------------------------------------------------------------
def algorithmic_bytes(b, ctx, hq, hkv, d, bs, esize=2):
    kv = b * 2 * ctx * hkv * d * esize
    io = b * hq * d * esize * 2 + b * ((ctx + bs - 1) // bs) * 4 + b * 4
    return kv, io


def peak_hbm():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, copy read+write)"
        except Exception:           # noqa: BLE001
            pass
    return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


def ncu_traffic(workload: str):
    """dram bytes per launch of the dominant kernel from the committed ncu capture, or None."""
    p = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)).get(workload)
        except Exception:           # noqa: BLE001
            return None
    return None


def make_workload(torch, dev, wl, layers, seed):
    """Synthetic KV pool [2, L, NB, Hkv, BS, D] (reference layout, paged_attention.py:26-30), fully
    scattered block tables (rows of a seeded randperm; no two sequences share a block)."""
    from vllm_from_scratch_b200 import paged_attention as P
    b, ctx, hq, hkv, d, bs = wl[:6]
    nblk = (ctx + bs - 1) // bs
    nb = b * nblk + 64
    pool = P.KVCachePool(nb, bs, hkv, d, layers, dtype=torch.bfloat16, device=dev)
    g = torch.Generator(device=dev).manual_seed(seed)
    flat = pool.kv_cache.view(-1)
    chunk = 1 << 28
    for lo in range(0, flat.numel(), chunk):        # fill in pieces: no 2x fp32 temporary of the pool
        n = min(chunk, flat.numel() - lo)
        flat[lo:lo + n] = torch.randn(n, generator=g, device=dev, dtype=torch.float32).to(torch.bfloat16)
    gc = torch.Generator().manual_seed(seed)
    perm = torch.randperm(nb, generator=gc)[: b * nblk].to(torch.int32).view(b, nblk)
    lens = torch.full((b,), ctx, dtype=torch.int32)
    q = torch.randn(b, hq, d, generator=gc).to(torch.bfloat16)
    k_new = torch.randn(b, hkv, d, generator=gc).to(torch.bfloat16)
    v_new = torch.randn(b, hkv, d, generator=gc).to(torch.bfloat16)
    pos = lens.long() - 1
    slots = perm.long()[torch.arange(b), pos // bs] * bs + pos % bs
    return pool, perm, lens, q, k_new, v_new, slots


def config_block(wl, b, hq, hkv, world, layers, kv_bytes, esize, head_mode=False, peer_mode=False, gather="peer"):
    """The `config` object of the JSON line.  The reference arm emits the SAME object for the same
    workload (it describes the workload, not the implementation), so the two arms compare like for like."""
    _, ctx, _, _, d, bs, _, desc = wl
    par = f"sequence-sharded x{world}, one pool per GPU, no collective"
    if head_mode:
        par = (f"KV-head-sharded x{world}: {hkv} KV / {hq} Q heads per GPU, " +
               ("output gather fused into the decode kernel (NVLink peer stores, " +
                ("fence + arrival counters" if gather == "peer-counter" else
                 "LL protocol: step number in every 8-byte word, no fence") + ", no collective call)"
                if peer_mode else f"one NCCL all-gather of [B, Hq/{world}, D] per step"))
    return {"workload": desc, "seqs_per_gpu": b, "context": ctx, "num_q_heads": hq, "num_kv_heads": hkv,
            "head_dim": d, "block_size": bs, "kv_dtype_bytes": esize,
            "kv_layout": "[2, L, NB, Hkv, BS, D] (reference pool)",
            "block_tables": "seeded randperm, fully scattered", "parallelism": par,
            "l2": f"inputs larger than L2: loop rotates over {layers} per-layer caches of "
                  f"{kv_bytes / 2**20:.0f} MiB each",
            "step": "one attention-layer decode call over the batch (bf16: ONE kernel - split-K decode that merges "
                    "its own partials; fp32: split-K kernel + reduce kernel)"}


class Timer:
    """K steps on the current stream between two CUDA events.  `lead` untimed steps are queued right before
    the first event WITHOUT draining the queue in between, so the timed steps start back to back behind work
    that is already running (no idle gap between a host-side synchronize and the first launch, which at
    --steps 20 used to cost 5 % of the region when 8 processes shared the host)."""

    def __init__(self, torch, dist, world):
        self.torch, self.dist, self.world = torch, dist, world
        self.e0 = torch.cuda.Event(enable_timing=True)
        self.e1 = torch.cuda.Event(enable_timing=True)

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()

    def run(self, step, steps, lead=2, first=0, many=None):
        """`many(first, n)` (optional) runs steps first .. first + n - 1 in as few launches as it can (e.g. one
        CUDA graph holding a whole round of layer caches); the step count is the same either way."""
        torch = self.torch
        self.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        if many is not None:
            many(first, lead)
        else:
            for i in range(lead):
                step(first + i)
        self.e0.record()
        if many is not None:
            many(first + lead, steps)
        else:
            for i in range(steps):
                step(first + lead + i)
        self.e1.record()
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        self.barrier()
        return self.e0.elapsed_time(self.e1), t0, t1

    def max_over_ranks(self, values, dev):
        if self.world == 1:
            return list(values)
        t = self.torch.tensor(list(values), device=dev, dtype=self.torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return t.tolist()


def capture_steps(torch, fn, n, one_graph=False):
    """One CUDA graph per i in range(n) of fn(i) (warm first: tensor maps, function attributes); with
    ``one_graph`` a single graph holding the n launches in order - what an engine's decode step over n layers
    looks like, and what lets consecutive launches overlap through programmatic dependent launch."""
    for i in range(n):
        fn(i)
    torch.cuda.synchronize()
    graphs = []
    if one_graph:
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for i in range(n):
                fn(i)
        return [g]
    for i in range(n):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            fn(i)
        graphs.append(g)
    return graphs


def strict_close(got, ref):
    """bf16 bound of tests/test_decode_strict.py: |got - ref| <= 2e-3 + 2^-7 |ref| elementwise."""
    return bool(((got.double() - ref.double()).abs() <= 2e-3 + 2.0 ** -7 * ref.double().abs()).all())


def head_sharded_section(torch, dist, P, S, dev, world, rank, steps, timer):
    """N > 1: the KV-head-sharded long-context workloads (BASELINE.json configs[2] and the 32k shape of
    configs[4]) with the fused NVLink gather and with NCCL, each parity-checked once before timing, next to the
    one-GPU time of the SAME batch measured in this run (every rank times it on its own GPU; max is reported)."""
    out = {}
    for name in ("c3", "c5h"):
        wl = WORKLOADS[name]
        b, ctx, hq, hkv, d, bs, layers, desc = wl
        if hkv % world:
            continue
        pool, bts_h, lens_h, q_h, _, _, _ = make_workload(torch, dev, wl, layers, seed=4321)       # same on every rank
        bts, lens, q = bts_h.to(dev), lens_h.to(dev), q_h.to(dev)
        kv_full_bytes, io_full = algorithmic_bytes(b, ctx, hq, hkv, d, bs)
        ws = P.new_decode_workspace(b, hq, hkv, d, dev)
        out_full = torch.empty_like(q)
        caches = [pool.layer(i) for i in range(layers)]
        # every arm is timed the same way: ONE CUDA graph holds a round over the layer caches (programmatic
        # dependent-launch edges kept, one host launch per round); `rounds` replays = rounds * layers steps
        rounds = max(steps // layers, 2)
        nsteps = rounds * layers
        graphs = capture_steps(torch, lambda i: P.paged_attention_decode_batched(q, caches[i], bts, lens, out=out_full, workspace=ws),
                               layers, one_graph=True)
        for i in range(3):
            graphs[0].replay()
        ms1, _, _ = timer.run(lambda i: graphs[0].replay(), rounds)
        ref_full = P.paged_attention_decode_batched(q, caches[0], bts, lens).clone()
        kvh, qh = S.shard_kv_heads(hkv, world, rank), S.shard_q_heads(hq, hkv, world, rank)
        local = pool.kv_cache[:, :, :, kvh.start:kvh.stop].contiguous()                             # [2, L, NB, Hkv / N, BS, D]
        del pool, caches, graphs
        torch.cuda.empty_cache()
        lcaches = [local[:, i] for i in range(layers)]
        hq_l, hkv_l = len(qh), len(kvh)
        q_l = q[:, qh.start:qh.stop].contiguous()
        ws_l = P.new_decode_workspace(b, hq_l, hkv_l, d, dev)
        plain = P.paged_attention_decode_batched(q_l, lcaches[0], bts, lens)
        want = torch.empty(world, b, hq_l, d, dtype=torch.bfloat16, device=dev)
        dist.all_gather_into_tensor(want.view(-1), plain.view(-1))
        entry = {"workload": desc, "batch": b, "context": ctx, "kv_heads_per_gpu": hkv_l, "layers_rotated": layers,
                 "one_gpu_us_per_step": None, "arms": {}}
        arms = {}
        # ---- fused gather (LL): decode + receive per layer cache as CUDA graphs, buffer sets alternate ----
        pg = S.PeerGather(b, hq, hkv, d, dev, world=world, rank=rank, protocol="ll")
        got = pg.decode(q_l, lcaches[0], bts, lens, workspace=ws_l, parity=0)
        pg.check()
        ok_bits = torch.equal(got, want)
        ok_full = strict_close(pg.full(got), ref_full)
        pg.decode(q_l, lcaches[1], bts, lens, workspace=ws_l, parity=1)                             # keep the parity sequence 0, 1, 0, 1 ...
        # a round = `layers` (even) steps: the 0 / 1 buffer-set alternation is preserved across replays
        pgraphs = capture_steps(torch, lambda i: pg.decode(q_l, lcaches[i], bts, lens, workspace=ws_l, parity=i % 2), layers,
                                one_graph=True)
        for i in range(3):
            pgraphs[0].replay()
        ms_p, _, _ = timer.run(lambda i: pgraphs[0].replay(), rounds)
        pg.check()
        arms["peer_ll"] = (ms_p / nsteps, ok_bits and ok_full,
                           "gather fused into the decode kernel: NVLink peer stores of {data, step} words, receive kernel; "
                           "decode + receive of a whole round over the layer caches replayed as one CUDA graph")
        pg.close()
        # ---- NCCL arm: the same decode kernel writes its slice in place + one all_gather_into_tensor ----
        hg = S.HeadGather(b, hq, d, torch.bfloat16, dev, world=world, rank=rank)

        def nccl_step(i):
            P.paged_attention_decode_batched(q_l, lcaches[i % layers], bts, lens, out=hg.local_out(), workspace=ws_l)
            hg.all_gather()

        nccl_step(0)
        torch.cuda.synchronize()
        ok_n = torch.equal(hg.buf, want) and strict_close(hg.full(), ref_full)
        how = "eager launches"
        try:
            ngraphs = capture_steps(torch, nccl_step, layers, one_graph=True)
            run_n = lambda i: ngraphs[0].replay()                                                   # noqa: E731
            how = "decode + all_gather of a whole round over the layer caches replayed as one CUDA graph"
        except Exception:                                                                           # noqa: BLE001 - capture of the collective refused: time it eagerly
            torch.cuda.synchronize()

            def run_n(i):
                for j in range(layers):
                    nccl_step(j)
        for i in range(3):
            run_n(i)
        ms_n, _, _ = timer.run(run_n, rounds)
        arms["nccl"] = (ms_n / nsteps, ok_n, "decode kernel + one NCCL all_gather_into_tensor of [B, Hq/N, D]; " + how)
        ms1_max, = timer.max_over_ranks([ms1 / nsteps], dev)
        entry["one_gpu_us_per_step"] = ms1_max * 1e3
        for aname, (ms, okv, what) in arms.items():
            ms_max, bad = timer.max_over_ranks([ms, 0.0 if okv else 1.0], dev)
            if bad:
                raise RuntimeError(f"bench.py head-sharded {name} / {aname}: gathered outputs differ from the ranks' own bits "
                                   "or from the full-head single-GPU decode - refusing to time a wrong result")
            entry["arms"][aname] = {"us_per_step": ms_max * 1e3, "tokens_per_s": b / (ms_max * 1e-3),
                                    "kv_read_gbs_per_gpu": kv_full_bytes / world / (ms_max * 1e-3) / 1e9,
                                    "pct_of_8tbs_per_gpu": 100.0 * kv_full_bytes / world / (ms_max * 1e-3) / 1e9 / NOMINAL_HBM_GBS,
                                    "speedup_vs_one_gpu": ms1_max / ms_max, "parity_checked": True, "what": what}
        out[name] = entry
        del local, lcaches
        torch.cuda.empty_cache()
    return out


def ramp_clocks(torch, step, ms=150.0, least=3):
    """Untimed: run step(i) for `ms` of wall clock (and at least `least` times).  Building a workload leaves the GPU
    idle long enough for the clocks to drop; a handful of warm-up steps does not bring them back."""
    t_end = time.perf_counter() + ms * 1e-3
    i = 0
    while i < least or time.perf_counter() < t_end:
        step(i)
        i += 1
        if i % 16 == 0:
            torch.cuda.synchronize()               # keep the launch queue bounded
    torch.cuda.synchronize()


def c5_point(torch, P, dev, world, rank, steps, timer):
    """BASELINE.json configs[4]: 512 seqs x 8k context over the box, sharded by sequence: 512 / N sequences per
    GPU (N = 1: the whole batch on one GPU, 16 GiB of KV per layer) - strong scaling on the stated batch."""
    if 512 % world:
        return None
    _, ctx, hq, hkv, d, bs, _, _ = WORKLOADS["c5"]
    b = 512 // world
    layers = 1 if b * ctx * hkv * d * 4 >= (8 << 30) else 2
    wl = (b, ctx, hq, hkv, d, bs, layers, "")
    pool, bts_h, lens_h, q_h, _, _, _ = make_workload(torch, dev, wl, layers, seed=777 + rank)
    bts, lens, q = bts_h.to(dev), lens_h.to(dev), q_h.to(dev)
    out = torch.empty_like(q)
    ws = P.new_decode_workspace(b, hq, hkv, d, dev)
    caches = [pool.layer(i) for i in range(layers)]
    graphs = capture_steps(torch, lambda i: P.paged_attention_decode_batched(q, caches[i], bts, lens, out=out, workspace=ws), layers)
    n = max(10, min(steps, 100))
    ramp_clocks(torch, lambda i: graphs[i % layers].replay())
    ms, _, _ = timer.run(lambda i: graphs[i % layers].replay(), n)
    ms_max, = timer.max_over_ranks([ms / n], dev)
    kv_bytes, _ = algorithmic_bytes(b, ctx, hq, hkv, d, bs)
    del pool, caches, graphs
    torch.cuda.empty_cache()
    return {"workload": "C5 whole-box decode: 512 seqs x 8192 ctx, GQA 32/8, head_dim 128, bf16, sharded by sequence",
            "seqs_total": 512, "seqs_per_gpu": b, "n_gpus": world, "scaling": "strong (total batch fixed at 512)",
            "steps": n, "ms_per_step": ms_max, "tokens_per_s": 512 / (ms_max * 1e-3),
            "kv_read_gbs_per_gpu": kv_bytes / (ms_max * 1e-3) / 1e9,
            "pct_of_8tbs_per_gpu": 100.0 * kv_bytes / (ms_max * 1e-3) / 1e9 / NOMINAL_HBM_GBS,
            "l2": f"{kv_bytes / 2**30:.0f} GiB of KV per layer cache, {layers} cache(s) rotated"}


def generic_f32_point(torch, P, dev, steps, timer, peak):
    """The generic streaming kernel (fp32 KV cache = the reference's default dtype, and every odd shape with
    16-byte aligned rows) on the C2 batch: KV bytes double.  Two launches per step (split-K + reduce)."""
    b, ctx, hq, hkv, d, bs, _, _ = WORKLOADS["c2"]
    layers = 2
    nblk = ctx // bs
    nb = b * nblk + 64
    pool = P.KVCachePool(nb, bs, hkv, d, layers, dtype=torch.float32, device=dev)
    g = torch.Generator(device=dev).manual_seed(5)
    flat = pool.kv_cache.view(-1)
    for lo in range(0, flat.numel(), 1 << 28):
        n = min(1 << 28, flat.numel() - lo)
        flat[lo:lo + n] = torch.randn(n, generator=g, device=dev, dtype=torch.float32)
    gc = torch.Generator().manual_seed(5)
    bts = torch.randperm(nb, generator=gc)[: b * nblk].to(torch.int32).view(b, nblk).to(dev)
    lens = torch.full((b,), ctx, dtype=torch.int32, device=dev)
    q = torch.randn(b, hq, d, generator=gc).to(dev)
    out = torch.empty_like(q)
    ws = P.new_decode_workspace(b, hq, hkv, d, dev)
    caches = [pool.layer(i) for i in range(layers)]
    graphs = capture_steps(torch, lambda i: P.paged_attention_decode_batched(q, caches[i], bts, lens, out=out, workspace=ws), layers)
    n = max(20, min(steps, 200))
    ramp_clocks(torch, lambda i: graphs[i % layers].replay(), least=4)
    ms, _, _ = timer.run(lambda i: graphs[i % layers].replay(), n)
    kv_bytes, io_bytes = algorithmic_bytes(b, ctx, hq, hkv, d, bs, esize=4)
    ach = (kv_bytes + io_bytes) / (ms / n * 1e-3) / 1e9
    del pool, caches, graphs
    torch.cuda.empty_cache()
    return {"workload": "C2 batch with an fp32 KV cache (64 x 2048, 32/8 heads, head_dim 128): generic kernel",
            "kernel": "pa_decode_stream_kernel<float> (cp.async-staged) + pa_reduce_kernel<float>", "launches_per_step": 2, "steps": n,
            "ms_per_step": ms / n, "tokens_per_s": b / (ms / n * 1e-3), "algorithmic_bytes_per_step": kv_bytes + io_bytes,
            "achieved_gbs": ach, "peak_gbs": peak, "frac": ach / peak, "frac_of_nominal_8tbs": ach / NOMINAL_HBM_GBS}


def run_ours(args):
    import torch
    import torch.distributed as dist
    from vllm_from_scratch_b200 import _lib, paged_attention as P

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py (impl=ours) needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        import datetime
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # NCCL prints its version banner on stdout when NCCL_DEBUG is set: keep stdout to the one JSON line
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=300))
    lib = _lib.load()
    timer = Timer(torch, dist, world)

    head_mode = args.shard == "head" and world > 1
    wl_name = args.workload or ("c3" if head_mode else "c2")
    wl = WORKLOADS[wl_name]
    b, ctx, hq, hkv, d, bs, def_layers, desc = wl
    layers = args.layers or def_layers
    gather = None
    if head_mode:
        # KV-head sharding: this rank holds Hkv / N heads of every block and the matching query heads;
        # block tables are identical on all ranks (same seed)
        from vllm_from_scratch_b200 import sharding as S
        hq_full = hq
        hkv, hq = len(S.shard_kv_heads(hkv, world, rank)), len(S.shard_q_heads(hq, hkv, world, rank))
        hkv_full = wl[3]
        wl = (b, ctx, hq, hkv, d, bs, def_layers, desc)
        if args.gather.startswith("peer"):
            # output gather fused into the decode kernel: rows go to every rank's buffer over NVLink
            # peer memory, arrival counters instead of a collective (sharding.PeerGather)
            gather = S.PeerGather(b, hq_full, hkv_full, d, dev, world=world, rank=rank,
                                  protocol="counter" if args.gather == "peer-counter" else "ll")
        else:
            gather = S.HeadGather(b, hq_full, d, torch.bfloat16, dev, world=world, rank=rank)
    pool, bts_h, lens_h, q_h, k_h, v_h, slots_h = make_workload(torch, dev, wl, layers,
                                                                seed=1234 + (0 if head_mode else rank))
    bts, lens, q = bts_h.to(dev), lens_h.to(dev), q_h.to(dev)
    peer_mode = head_mode and args.gather.startswith("peer")
    out = gather.local_out() if head_mode else torch.empty_like(q)
    ws = P.new_decode_workspace(b, hq, hkv, d, dev)
    caches = [pool.layer(i) for i in range(layers)]
    kv_bytes, io_bytes = algorithmic_bytes(b, ctx, hq, hkv, d, bs)

    def step_eager(i):
        if peer_mode:
            gather.decode(q, caches[i % layers], bts, lens, workspace=ws, parity=i % 2)   # decode + fused gather + receive
            return
        P.paged_attention_decode_batched(q, caches[i % layers], bts, lens, out=out, workspace=ws)
        if head_mode:
            gather.all_gather()             # the kernel wrote its slice of the gather buffer in place

    # The step is replayed from one CUDA graph per layer cache: the kernel (93 us at C2, 47 us head-sharded at
    # 8 GPUs) is shorter than the Python -> ctypes launch path when 8 processes share the host.  The graph
    # holds exactly what the eager step launches (one kernel; + the receive kernel with the fused gather).
    if peer_mode and layers % 2:
        raise SystemExit("--gather peer needs an even --layers (the two gather buffer sets alternate)")
    graphed = not (head_mode and not peer_mode)          # the NCCL arm stays eager (collective outside a graph)
    step_many = None
    if graphed:
        if peer_mode:
            for i in range(2 * layers):                  # eager first, an even count: both buffer sets warm
                step_eager(i)
        graphs = capture_steps(torch, step_eager, layers)
        step = lambda i: graphs[i % layers].replay()     # noqa: E731
        if not head_mode:
            # one more graph holding a whole round over the layer caches, in order: what an engine's decode step
            # over its layers looks like - consecutive launches inside one graph keep their programmatic
            # dependent-launch edges (the next kernel's prologue runs under the previous kernel's tail), which
            # separate graph launches cannot.  The timed region still counts every kernel launch as one step;
            # steps that do not fill a round are replayed from the per-layer graphs.
            round_graph = capture_steps(torch, step_eager, layers, one_graph=True)[0]

            def step_many(first, n):
                i = first
                while n > 0:
                    if i % layers == 0 and n >= layers:
                        round_graph.replay()
                        i, n = i + layers, n - layers
                    else:
                        graphs[i % layers].replay()
                        i, n = i + 1, n - 1
    else:
        step = step_eager

    sampler = ClockSampler(local)
    sampler.start()
    # untimed: ramp clocks, fault in the tensor maps, then the W warm-up steps
    t_end = time.perf_counter() + args.ramp_ms * 1e-3
    i = 0
    # head mode: every rank must run the SAME number of steps (collective / arrival counts)
    while (i < int(args.ramp_ms * 4) - int(args.ramp_ms * 4) % (2 * layers)) if head_mode else (time.perf_counter() < t_end):
        step(i)
        i += 1
        if i % 64 == 0:
            torch.cuda.synchronize()
    warm = max(args.warmup, 3)
    warm += (-warm) % (2 * layers) if peer_mode else 0   # keep the 0 / 1 buffer-set alternation aligned
    for i in range(warm):
        step(i)
    torch.cuda.synchronize()

    # ---- timed region: K steps, device time, max over ranks --------------------------------
    lead = 2 * layers if peer_mode else (layers if step_many is not None else 2)
    ms_total, t0, t1 = timer.run(step, args.steps, lead=lead, many=step_many)
    clocks = sampler.summary(t0, t1)
    if peer_mode and (args.steps % 2):                   # odd K: one more step realigns the buffer sets
        step(lead + args.steps)

    # ---- per-kernel durations with events around each launch (eager; cross-check of the roofline) ----
    nprof = min(args.steps, 512)
    nprof -= nprof % 2 if peer_mode else 0
    dec_avg = red_avg = None
    got = 0
    if nprof > 0:
        lib.pa_profile_begin(nprof)
        for i in range(nprof):
            step_eager(i)
        import ctypes
        dec_ms = (ctypes.c_float * nprof)()
        red_ms = (ctypes.c_float * nprof)()
        got = lib.pa_profile_end(dec_ms, red_ms, nprof)
        if got <= 0:
            raise RuntimeError(f"pa_profile_end returned {got}")
        dec_avg = sum(dec_ms[:got]) / got
        red_avg = sum(red_ms[:got]) / got

    # ---- e2e: host buffers in, host buffer out, through the public Python API ----------------
    # graph.DecodeStep: the step's inputs (q, new K / V rows, their slots, block tables, sequence
    # lengths) sit in ONE pinned staging buffer and cross PCIe as ONE copy; H2D -> reshape_and_cache
    # -> decode -> D2H(out) is one captured CUDA graph per layer cache; the host waits for the result
    # every step.
    from vllm_from_scratch_b200.graph import DecodeStep
    ne2e = args.e2e_steps or min(args.steps, 300)
    steps_e2e = []
    for c in caches:
        st = DecodeStep(c, b, hq, bts_h.shape[1], write_kv=True, out=out if head_mode else None)
        st.host["q"].copy_(q_h)
        st.host["k"].copy_(k_h)
        st.host["v"].copy_(v_h)
        st.host["slots"].copy_(slots_h)
        st.host["block_tables"].copy_(bts_h)
        st.host["seq_lens"].copy_(lens_h)
        if not head_mode:
            st.capture()
        steps_e2e.append(st)
    h2d, d2h = steps_e2e[0].h2d_bytes, steps_e2e[0].d2h_bytes

    if peer_mode:
        # host buffers in, gathered host buffer out, through sharding.PeerGather: one staged H2D copy of
        # q + new K / V rows (+ slots), slot write, decode with the fused gather, D2H of ALL ranks' rows
        pin = lambda t: t.contiguous().pin_memory()                                     # noqa: E731
        hq_p, hk_p, hv_p, hs_p = pin(q_h), pin(k_h), pin(v_h), pin(slots_h)
        k_d, v_d, s_d = k_h.to(dev), v_h.to(dev), slots_h.to(dev)
        full_h = torch.empty((world, b, hq, d), dtype=torch.bfloat16).pin_memory()
        h2d = sum(t.numel() * t.element_size() for t in (hq_p, hk_p, hv_p, hs_p))
        d2h = full_h.numel() * full_h.element_size()

    def e2e_step_peer(i):
        q.copy_(hq_p, non_blocking=True)
        k_d.copy_(hk_p, non_blocking=True)
        v_d.copy_(hv_p, non_blocking=True)
        s_d.copy_(hs_p, non_blocking=True)
        P.reshape_and_cache(k_d, v_d, caches[i % layers], s_d)
        full_h.copy_(gather.decode(q, caches[i % layers], bts, lens, workspace=ws, parity=i % 2), non_blocking=True)
        torch.cuda.current_stream().synchronize()

    def e2e_step(i):
        if peer_mode:
            return e2e_step_peer(i)
        st = steps_e2e[i % layers]
        st.replay()
        if head_mode:
            gather.all_gather()
        st.wait()                                                     # the caller reads st.out now
        if head_mode:
            torch.cuda.current_stream().synchronize()

    ne2e -= ne2e % 2 if peer_mode else 0
    # Untimed ramp of this leg: building and capturing the steps above left the GPU idle for tens of ms, and a
    # step here keeps it only ~75 % busy, so the clocks take a few hundred ms of steps to come back
    # (tools/exp_e2e_jitter.py: 177 us per step for the first ~0.3 s, 116 us from then on - every step, not
    # outliers).  Six steps to size the ramp, then --ramp-ms worth of steps; the count is agreed over ranks
    # (the multi-GPU legs hold collectives / peer hand-shakes).
    t_w = time.perf_counter()
    for i in range(6):
        e2e_step(i)
    torch.cuda.synchronize()
    per = max((time.perf_counter() - t_w) / 6, 1e-6)
    n_ramp = int(timer.max_over_ranks([min(args.ramp_ms * 1e-3 / per, 10000.0)], dev)[0])
    n_ramp += n_ramp % 2
    for i in range(n_ramp):
        e2e_step(i)
    ms_e2e, _, _ = timer.run(e2e_step, ne2e, lead=0)
    sampler.stop_flag.set()

    ms_total, ms_e2e, dec_avg_m, red_avg_m = timer.max_over_ranks([ms_total, ms_e2e, dec_avg or 0.0, red_avg or 0.0], dev)
    ms_step = ms_total / args.steps
    seqs = b if head_mode else world * b                  # head mode: all ranks work on the SAME batch
    value = seqs / (ms_step * 1e-3)
    e2e_value = seqs / (ms_e2e / ne2e * 1e-3)
    kv_gbs = world * kv_bytes / (ms_step * 1e-3) / 1e9
    peak, peak_src = peak_hbm()
    # The timed region holds ONE decode kernel per step and nothing else on the stream (graph replays back to
    # back), so region / launches IS the kernel's average duration including its launch gap - an upper bound.
    kernel_ms = ms_step
    achieved = (kv_bytes + io_bytes) / (kernel_ms * 1e-3) / 1e9

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "strong" if head_mode else "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": config_block(wl, b, hq, hkv, world, layers, kv_bytes, 2, head_mode, peer_mode, args.gather),
        "kv_read_gbs": kv_gbs, "kv_read_gbs_per_gpu": kv_gbs / world,
        "pct_of_8tbs_per_gpu": 100.0 * kv_gbs / world / NOMINAL_HBM_GBS,
        "kv_bytes_per_step_per_gpu": kv_bytes,
        "launch_path": (("one CUDA graph per round over the layer caches (the eager steps' launches in order, programmatic "
                         "dependent-launch edges kept), replayed; steps that do not fill a round from one graph per layer cache"
                         if step_many is not None else
                         "one CUDA graph per layer cache, replayed (captures exactly the eager step's launches)")
                        if graphed else "eager launches"),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": ncu_traffic(wl_name), "peak_source": peak_src,
                     "kernel": "pa_decode_ws_kernel<128>", "kernel_ms": kernel_ms,
                     "kernel_ms_source": "CUDA events over the timed region / launches (one kernel per step, max over ranks)",
                     "kernel_ms_event_pairs": dec_avg_m if got else None,
                     "kernel_ms_event_pairs_note": "the same kernel launched eagerly with an event pair around every launch "
                                                   "(adds ~2 us of event overhead per launch); cross-check only",
                     "algorithmic_bytes_per_launch": kv_bytes + io_bytes,
                     "frac_of_nominal_8tbs": achieved / NOMINAL_HBM_GBS, "launches_timed": args.steps},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "steps": ne2e, "warmup": 6 + n_ramp, "ms_per_step": ms_e2e / ne2e,
                "what": "graph.DecodeStep: every input of the step starts in pinned host memory; new K / V rows, "
                        "slots, block tables and seq lens cross PCIe as one staged H2D copy, the decode kernel reads "
                        "q from and writes out to pinned host memory in place (zero-copy over PCIe; both counted in "
                        "the byte totals); copy -> reshape_and_cache -> decode is one CUDA graph and the host waits "
                        "for the result every step; the KV pool is resident state"},
        "gpu_launches": args.steps * (2 if peer_mode else 1),   # timed region: one decode kernel per step (+ the receive kernel with the fused gather)
        "clocks": clocks,
    }
    # ---- the other driver-visible points ------------------------------------------------------------------
    want_cpu = rank == 0 and world == 1 and not args.no_cpu_baseline
    cpu_kv = pool.kv_cache[:, 0].cpu() if want_cpu else None
    del pool, caches, steps_e2e
    if graphed:
        del graphs
    torch.cuda.empty_cache()
    default_run = args.workload is None and not head_mode and not args.no_extras
    if default_run:
        xs = max(args.steps, 50)
        line["c5_512x8k"] = c5_point(torch, P, dev, world, rank, xs, timer)
        if world == 1:
            line["generic_f32"] = generic_f32_point(torch, P, dev, xs, timer, peak)
        else:
            from vllm_from_scratch_b200 import sharding as S
            line["head_sharded"] = head_sharded_section(torch, dist, P, S, dev, world, rank, max(args.steps, 100), timer)
    if want_cpu:
        line["cpu_baseline"] = cpu_baseline(torch, cpu_kv, bts_h, lens_h, q_h, wl, args.cpu_seconds)
    if rank == 0:
        emit(line)
    if world > 1:
        dist.destroy_process_group()


# ---------------------------------------------------------------------------------------
# CPU legs: the oracle port of the reference algorithm (oracle/ is the checker / baseline only)
# ---------------------------------------------------------------------------------------
def cpu_model_name():
    try:
        for ln in open("/proc/cpuinfo"):
            if ln.startswith("model name"):
                return ln.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def cpu_baseline(torch, kv_layer_cpu, bts, lens, q, wl, budget_s):
    """Reference algorithm as shipped (row-by-row gather through the block table, then einsum /
    softmax / einsum in bf16: oracle decode_loop = paged_attention.py:80-101) on ONE host thread
    (it is interpreter-bound), on a bounded sample of the same batch; plus the vectorised
    restatement on all cores for a fairer CPU number."""
    from oracle import paged_attention_oracle as O
    b, ctx, hq, hkv, d, bs = wl[:6]
    t0 = time.perf_counter()
    done = 0
    while done < b and (time.perf_counter() - t0 < budget_s or done == 0):
        O.decode_loop(q[done], kv_layer_cpu, bts[done].tolist(), int(lens[done]), d, hq, hkv, bs)
        done += 1
    dt = time.perf_counter() - t0
    tv0 = time.perf_counter()
    nvec = 0
    while nvec < b and (time.perf_counter() - tv0 < 3.0 or nvec == 0):
        O.decode_vectorised(q[nvec], kv_layer_cpu, bts[nvec].tolist(), int(lens[nvec]), d, hq, hkv, bs,
                            compute_dtype=torch.float32)
        nvec += 1
    dtv = time.perf_counter() - tv0
    return {"value": done / dt, "unit": UNIT, "cores": 1, "kind": "port",
            "sample": f"{done} of {b} sequences of the same batch (ctx {ctx}), oracle decode_loop, bf16, "
                      f"{dt:.1f} s; sequences are independent so tokens/s = sequences / time",
            "cpu_model": cpu_model_name(), "host_cores": os.cpu_count(),
            "vectorised_port": {"value": nvec / dtv, "unit": UNIT, "cores": torch.get_num_threads(),
                                "sample": f"{nvec} sequences, oracle decode_vectorised fp32, {dtv:.1f} s"}}


_REF_STATE = {}


def _ref_worker(idx):
    from oracle import paged_attention_oracle as O
    s = _REF_STATE
    b, ctx, hq, hkv, d, bs = s["wl"][:6]
    O.decode_loop(s["q"][idx], s["kv"], s["bts"][idx].tolist(), ctx, d, hq, hkv, bs)
    return idx


def run_reference(args):
    """--impl reference: the reference's own CPU algorithm (oracle port decode_loop = paged_attention.py:80-101,
    bit-identical to the reference function) on all host cores.  A step is the SAME batch as the GPU arm's
    (all `b` sequences of the workload, sequences are independent: one per worker process, in waves of
    `cores`); only the step COUNT is capped so that the run ends within a few minutes."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import multiprocessing as mp

    import torch
    torch.set_num_threads(1)
    wl = WORKLOADS[args.workload or ("c3" if args.shard == "head" and args.gpus > 1 else "c2")]
    b, ctx, hq, hkv, d, bs, def_layers, desc = wl
    cores = max(1, min(os.cpu_count() or 1, b, 64))
    nblk = (ctx + bs - 1) // bs
    nb = b * nblk + 64
    g = torch.Generator().manual_seed(1234)
    kv = torch.randn(2, nb, hkv, bs, d, generator=g).to(torch.bfloat16)
    bts = torch.randperm(nb, generator=g)[: b * nblk].to(torch.int32).view(b, nblk)
    q = torch.randn(b, hq, d, generator=g).to(torch.bfloat16)
    _REF_STATE.update(wl=wl, kv=kv, bts=bts, q=q)
    ctxmp = mp.get_context("fork")
    with ctxmp.Pool(cores) as pool:
        t0 = time.perf_counter()
        pool.map(_ref_worker, range(b), chunksize=1)               # warm-up step 1, also the step-time estimate
        est = time.perf_counter() - t0
        budget = float(os.environ.get("BENCH_REF_SECONDS", "150"))
        steps = max(1, min(args.steps, int(budget * 0.8 / est)))
        warmup = max(1, min(args.warmup, int(budget * 0.2 / est)))
        for _ in range(warmup - 1):
            pool.map(_ref_worker, range(b), chunksize=1)
        t0 = time.perf_counter()
        for _ in range(steps):
            pool.map(_ref_worker, range(b), chunksize=1)
        dt = time.perf_counter() - t0
    value = b * steps / dt
    kv_bytes, _ = algorithmic_bytes(b, ctx, hq, hkv, d, bs)
    sample = (f"each step = all {b} sequences of the batch (ctx {ctx}), one per worker process in waves of {cores} on "
              f"{cores} host cores, oracle decode_loop in bf16; {steps} timed steps"
              + (f" (capped from {args.steps} to bound the run)" if steps != args.steps else ""))
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": warmup, "ms_per_step": dt / steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": config_block(wl, b, hq, hkv, max(args.gpus, 1), args.layers or def_layers, kv_bytes, 2),
        "reference_note": "CPU arm: runs on rank 0's host cores only; one layer cache (nothing on a CPU is L2-resident "
                          "at 512 MiB); `config` is the GPU arm's workload description",
        "kv_read_gbs": kv_bytes * steps / dt / 1e9,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample,
                         "cpu_model": cpu_model_name(), "host_cores": os.cpu_count()},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


_JSON_FD = None


def emit(line) -> None:
    """The ONE JSON line goes to the process's original stdout."""
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    global _JSON_FD
    args = parse_args()
    # Libraries write to fd 1 behind Python's back (NCCL prints its version banner there on
    # communicator creation): point fd 1 at stderr for the whole run and keep the real stdout
    # for the one JSON line.
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    if os.environ.get("BENCH_DEBUG"):
        import faulthandler
        faulthandler.dump_traceback_later(int(os.environ.get("BENCH_DEBUG_AFTER", "60")), exit=True)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
