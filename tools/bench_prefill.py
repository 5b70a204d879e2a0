"""Throughput of the tcgen05 paged prefill kernel against the tensor
roofline.  One JSON line per shape: TFLOP/s = 4 * Hq * D * sum_q (pos_q + 1) / kernel time (CUDA
events around the launch on its stream, median of `--iters` after warm-up), peak = MEASURED_PEAKS
bf16 TF/s.  The KV pool of every shape is larger than it needs to be reused from L2 only by the
CTAs that share a KV head - which is the point of the GQA-packed tile.

    python tools/bench_prefill.py [--iters 20] [--shape full8k|batch2k|chunk32k|mixed]
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vllm_from_scratch_b200 import paged_attention as P            # noqa: E402

HQ, HKV, D, BS = 32, 8, 128, 16
# name -> list of (ctx_len, q_len)
SHAPES = {
    "full8k": [(8192, 8192)],
    "batch2k": [(2048, 2048)] * 8,
    "chunk32k": [(32768, 2048)],
    "mixed": [(2048, 512)] * 4 + [(2048, 1)] * 60,
    "mixed_split": [(2048, 512)] * 4 + [(2048, 1)] * 60,
    "full32k": [(32768, 32768)],
}


def peak():
    try:
        m = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(m["bf16_tflops"]), "measured cuBLAS bf16 burst (MEASURED_PEAKS.json)"
    except Exception:                                               # noqa: BLE001
        return 1500.0, "fallback"


def run(name, iters, warmup):
    specs = SHAPES[name]
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(0)
    nblk = [(c + BS - 1) // BS for c, _ in specs]
    nb = sum(nblk) + 16
    kv = torch.randn(2, nb, HKV, BS, D, generator=g, device=dev, dtype=torch.float32).to(torch.bfloat16)
    perm = torch.randperm(nb, generator=torch.Generator().manual_seed(0)).to(torch.int32)
    bts = torch.zeros(len(specs), max(nblk), dtype=torch.int32)
    at = 0
    for i, n in enumerate(nblk):
        bts[i, :n] = perm[at:at + n]
        at += n
    cu = torch.tensor([0] + list(torch.tensor([q for _, q in specs]).cumsum(0)), dtype=torch.int32)
    ctx = torch.tensor([c for c, _ in specs], dtype=torch.int32)
    total_q = int(cu[-1])
    q = torch.randn(total_q, HQ, D, generator=g, device=dev, dtype=torch.float32).to(torch.bfloat16)
    out = torch.empty_like(q)
    bts_d, cu_d, ctx_d = bts.to(dev), cu.to(dev), ctx.to(dev)
    maxq = max(ql for _, ql in specs)
    flops = sum(4 * HQ * D * sum(range(c - ql + 1, c + 1)) for c, ql in specs)

    split = name.endswith("_split")
    if split:
        # the adapter's route for decode-heavy mixed steps: the varlen launch skips the one-token
        # requests (ctx 0), the decode kernel takes them
        is_dec = torch.tensor([ql == 1 for _, ql in specs])
        ctx_pf = torch.where(is_dec, torch.zeros_like(ctx), ctx).to(dev)
        rows = cu[:-1][is_dec].long().to(dev)
        bts_dec, ctx_dec = bts_d[is_dec.to(dev)].contiguous(), ctx_d[is_dec.to(dev)].contiguous()

    def step():
        if split:
            P.paged_attention_prefill_varlen(q, kv, bts_d, cu_d, ctx_pf, maxq, out=out)
            out.index_copy_(0, rows, P.paged_attention_decode_batched(q.index_select(0, rows), kv, bts_dec, ctx_dec))
            return
        P.paged_attention_prefill_varlen(q, kv, bts_d, cu_d, ctx_d, maxq, out=out)

    for _ in range(warmup):
        step()
    torch.cuda.synchronize()
    times = []
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        step()
        e1.record()
        e1.synchronize()
        times.append(e0.elapsed_time(e1))
    times.sort()
    ms = times[len(times) // 2]
    pk, src = peak()
    tf = flops / (ms * 1e-3) / 1e12
    print(json.dumps({"kernel": "pa_prefill_tc_kernel<128>", "shape": name, "specs": f"{len(specs)} seqs, (ctx, q_len) = {specs[0]}" + (" ..." if len(specs) > 1 else ""),
                      "heads": f"{HQ}/{HKV}", "head_dim": D, "q_tokens": total_q, "ms": ms, "ms_min": times[0],
                      "tflops": tf, "roofline": {"bound": "tensor", "achieved": tf, "peak": pk, "unit": "TFLOP/s",
                                                 "frac": tf / pk, "peak_source": src},
                      "tokens_per_s": total_q / (ms * 1e-3), "iters": iters}))


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--shape", default=None)
    a = ap.parse_args()
    for n in ([a.shape] if a.shape else ["full8k", "batch2k", "chunk32k", "mixed", "mixed_split"]):
        run(n, a.iters, a.warmup)
