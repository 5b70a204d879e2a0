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
        dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=180))
    lib = _lib.load()

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

    def step(i):
        if peer_mode:
            gather.decode(q, caches[i % layers], bts, lens, workspace=ws, parity=i % 2)   # decode + fused gather + wait
            return
        P.paged_attention_decode_batched(q, caches[i % layers], bts, lens, out=out, workspace=ws)
        if head_mode:
            gather.all_gather()             # the kernel wrote its slice of the gather buffer in place

    step_eager = step
    if peer_mode:
        # decode + fused gather + arrival wait per layer cache as CUDA graphs (the wait's step number
        # lives on the device): at 8 GPUs the kernel is shorter than the host launch path
        if layers % 2:
            raise SystemExit("--gather peer needs an even --layers (the two gather buffer sets alternate)")
        for i in range(2 * layers):                        # eager first: tensor maps, workspaces; even count
            step_eager(i)
        torch.cuda.synchronize()
        graphs = []
        for i in range(layers):
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr):
                gather.decode(q, caches[i], bts, lens, workspace=ws, parity=i % 2)
            graphs.append(gr)

        def step(i):                                       # noqa: F811
            graphs[i % layers].replay()

    def barrier():
        if world > 1:
            dist.barrier()

    sampler = ClockSampler(local)
    sampler.start()
    # untimed: ramp clocks, fault in the tensor maps, then the W warm-up steps
    t_end = time.perf_counter() + args.ramp_ms * 1e-3
    i = 0
    # head mode: every rank must run the SAME number of steps (collective / arrival counts)
    while (i < int(args.ramp_ms * 4)) if head_mode else (time.perf_counter() < t_end):
        step(i)
        i += 1
        if i % 64 == 0:
            torch.cuda.synchronize()
    for i in range(max(args.warmup, 3)):
        step(i)
    torch.cuda.synchronize()

    # ---- timed region: K steps, device time, max over ranks --------------------------------
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    e0.record()
    for i in range(args.steps):
        step(i)
    e1.record()
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    barrier()
    ms_total = e0.elapsed_time(e1)
    clocks = sampler.summary(t0, t1)

    # ---- per-kernel durations (roofline): same steps with events around each kernel -----------
    nprof = min(args.steps, 512)
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

    for i in range(5):
        e2e_step(i)
    barrier()
    torch.cuda.synchronize()
    e0.record()
    for i in range(ne2e):
        e2e_step(i)
    e1.record()
    torch.cuda.synchronize()
    barrier()
    ms_e2e = e0.elapsed_time(e1)
    sampler.stop_flag.set()

    if world > 1:
        t = torch.tensor([ms_total, ms_e2e, dec_avg, red_avg], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total, ms_e2e, dec_avg, red_avg = t.tolist()
    ms_step = ms_total / args.steps
    seqs = b if head_mode else world * b                  # head mode: all ranks work on the SAME batch
    value = seqs / (ms_step * 1e-3)
    e2e_value = seqs / (ms_e2e / ne2e * 1e-3)
    kv_gbs = world * kv_bytes / (ms_step * 1e-3) / 1e9
    peak, peak_src = peak_hbm()
    achieved = (kv_bytes + io_bytes) / (dec_avg * 1e-3) / 1e9

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "strong" if head_mode else "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": desc, "seqs_per_gpu": b, "context": ctx, "num_q_heads": hq, "num_kv_heads": hkv,
                   "head_dim": d, "block_size": bs, "kv_layout": "[2, L, NB, Hkv, BS, D] (reference pool)",
                   "block_tables": "seeded randperm, fully scattered",
                   "parallelism": (f"KV-head-sharded x{world}: {hkv} KV / {hq} Q heads per GPU, " +
                                   ("output gather fused into the decode kernel (NVLink peer stores, " +
                                    ("fence + arrival counters" if args.gather == "peer-counter" else
                                     "LL protocol: step number in every 8-byte word, no fence") +
                                    ", no collective call)" if peer_mode else
                                    f"one NCCL all-gather of [B, Hq/{world}, D] per step") if head_mode else
                                   f"sequence-sharded x{world}, one pool per GPU, no collective"),
                   "l2": f"inputs larger than L2: loop rotates over {layers} per-layer caches of "
                         f"{kv_bytes / 2**20:.0f} MiB each",
                   "step": "one attention-layer decode call (one fused split-K + merge kernel) over the batch"},
        "kv_read_gbs": kv_gbs, "kv_read_gbs_per_gpu": kv_gbs / world,
        "pct_of_8tbs_per_gpu": 100.0 * kv_gbs / world / NOMINAL_HBM_GBS,
        "kv_bytes_per_step_per_gpu": kv_bytes,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": ncu_traffic(wl_name), "peak_source": peak_src,
                     "kernel": "pa_decode_ws_kernel<128>", "kernel_ms": dec_avg, "reduce_kernel_ms": red_avg,
                     "algorithmic_bytes_per_launch": kv_bytes + io_bytes,
                     "frac_of_nominal_8tbs": achieved / NOMINAL_HBM_GBS, "launches_timed": got},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "steps": ne2e, "ms_per_step": ms_e2e / ne2e,
                "what": "graph.DecodeStep: every input of the step starts in pinned host memory; new K / V rows, "
                        "slots, block tables and seq lens cross PCIe as one staged H2D copy, the decode kernel reads "
                        "q from and writes out to pinned host memory in place (zero-copy over PCIe; both counted in "
                        "the byte totals); copy -> reshape_and_cache -> decode is one CUDA graph and the host waits "
                        "for the result every step; the KV pool is resident state"},
        "gpu_launches": args.steps,       # timed region: one decode launch per step (the fast path has no reduce launch)
        "clocks": clocks,
    }
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(torch, pool.kv_cache[:, 0].cpu(), bts_h, lens_h, q_h, wl, args.cpu_seconds)
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
    """--impl reference: the reference's own CPU algorithm (oracle port, decode_loop) on all host
    cores - one process per core, one sequence per process per step (sequences are independent)."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    import multiprocessing as mp

    import torch
    torch.set_num_threads(1)
    wl = WORKLOADS[args.workload or ("c3" if args.shard == "head" and args.gpus > 1 else "c2")]
    b, ctx, hq, hkv, d, bs, _, desc = wl
    cores = max(1, min(os.cpu_count() or 1, b, 64))
    nseq = cores                                   # sample: one sequence per core per step
    nblk = (ctx + bs - 1) // bs
    nb = nseq * nblk + 8
    g = torch.Generator().manual_seed(1234)
    kv = torch.randn(2, nb, hkv, bs, d, generator=g).to(torch.bfloat16)
    bts = torch.randperm(nb, generator=g)[: nseq * nblk].to(torch.int32).view(nseq, nblk)
    q = torch.randn(nseq, hq, d, generator=g).to(torch.bfloat16)
    _REF_STATE.update(wl=wl, kv=kv, bts=bts, q=q)
    # keep the whole run within a few minutes: ~0.5-1.5 s per step
    steps = min(args.steps, 20)
    warmup = min(max(args.warmup, 1), 2)
    ctxmp = mp.get_context("fork")
    with ctxmp.Pool(cores) as pool:
        for _ in range(warmup):
            pool.map(_ref_worker, range(nseq), chunksize=1)
        t0 = time.perf_counter()
        for _ in range(steps):
            pool.map(_ref_worker, range(nseq), chunksize=1)
        dt = time.perf_counter() - t0
    value = nseq * steps / dt
    kv_bytes, _ = algorithmic_bytes(nseq, ctx, hq, hkv, d, bs)
    sample = (f"each step = {nseq} of the batch's {b} sequences (ctx {ctx}), one per process on {cores} cores, "
              f"oracle decode_loop in bf16; {steps} steps (capped from {args.steps} to bound the run)")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": warmup, "ms_per_step": dt / steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": desc, "seqs_per_gpu": b, "context": ctx, "num_q_heads": hq, "num_kv_heads": hkv,
                   "head_dim": d, "block_size": bs, "note": "CPU arm: runs on rank 0's host cores only"},
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
