"""Config C4 (BASELINE.json configs[3]): the engine loop - mixed chunked-prefill + decode batches with
prefix-cache block sharing and swap-in / swap-out - through engine_adapter + kv_manager.

    python tools/bench_engine.py [--reference]

The workload is the reference's own ``benchmark_throughput`` (05_mini_vllm/mini_vllm.py:935-975: vocab
1000, hidden 256, 4 layers, 8 heads, inter 1024, block_size 16, max_num_seqs 16; 16 prompts of 10-50
tokens, 20 new tokens each), with SURVEY.md 8(d)'s additions: 8 of the 16 prompts share a 32-token
(2-block) prefix, the token budget is 64 (forces chunking) and the pool is small enough (40 blocks) that
requests are preempted by swap.  Toy sizes are launch-bound, so this is reported as engine tok/s, not as
a roofline point.  One JSON line per engine:

  * ``tiny_cuda``  - the TinyEngine harness (tests/engine_harness.py; /root/reference does not exist on
    the GPU box) on cuda through the CUDA kernels, fp32 (D 32, generic path) and bf16 hidden 1024
    (D 128: one tcgen05 varlen launch per step);
  * ``tiny_cpu_oracle`` - the same engine on CPU through the oracle (the CPU baseline of this config);
  * ``--reference`` (build container only): the unmodified reference engine on CPU, as shipped (its
    approximate attention, no sharing, no swap: with the 40-block pool it would stall, so it gets 256).
"""
import argparse
import json
import os
import random
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from engine_harness import OracleOps, TinyEngine  # noqa: E402
from vllm_from_scratch_b200 import engine_adapter, kv_manager  # noqa: E402


def prompts():
    rng = random.Random(0)
    prefix = [rng.randint(1, 999) for _ in range(32)]
    out = []
    for i in range(16):
        body = [rng.randint(1, 999) for _ in range(rng.randint(10, 50))]
        out.append(prefix + body[:18] if i % 2 == 0 else body)
    return out


def run_tiny(device, ops, dtype, hidden, heads, blocks, label, graphs=None):
    eng = TinyEngine(vocab=1000, hidden=hidden, layers=4, heads=heads, inter=4 * hidden, num_blocks=blocks, block_size=16,
                     max_seqs=16, token_budget=64, device=device, dtype=dtype, hashed_blocks=True)
    engine_adapter.install(eng, ops=ops, cuda_graphs=graphs)
    mgr = kv_manager.install(eng, num_host_blocks=256, ops=ops)
    ps = prompts()

    def once():
        reqs = [eng.add_request(p, max_new_tokens=20) for p in ps]
        steps = eng.run()
        return reqs, steps

    once()                                                            # warm-up: tensor maps, workspaces, allocator caches
    if device != "cpu":
        torch.cuda.synchronize()
    before = (mgr.stats.prefix_hit_tokens, mgr.stats.preemptions, mgr.stats.blocks_swapped_out)
    t0 = time.perf_counter()
    reqs, steps = once()
    if device != "cpu":
        torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    out_tok = sum(len(r.output_token_ids) for r in reqs)
    in_tok = sum(len(p) for p in ps)
    print(json.dumps({"engine": label, "device": device, "dtype": str(dtype).replace("torch.", ""), "hidden": hidden,
                      "heads": heads, "head_dim": hidden // heads, "kv_blocks": blocks, "requests": len(ps),
                      "input_tokens": in_tok, "output_tokens": out_tok, "steps": steps, "seconds": dt,
                      "output_tok_per_s": out_tok / dt, "total_tok_per_s": (in_tok + out_tok) / dt,
                      "decode_graphs": None if eng._pa_b200["graphs"] is None else
                      {"captures": eng._pa_b200["graphs"].captures, "replays": eng._pa_b200["graphs"].replays},
                      "table_entries_shipped": eng._pa_b200["tables"].updates_sent,
                      "prefix_hit_tokens": mgr.stats.prefix_hit_tokens - before[0],
                      "preemptions": mgr.stats.preemptions - before[1],
                      "blocks_swapped_out": mgr.stats.blocks_swapped_out - before[2]}), flush=True)


def run_reference():
    sys.path.insert(0, "/root/reference/05_mini_vllm")
    import mini_vllm
    torch.manual_seed(0)
    eng = mini_vllm.MiniVLLM(vocab_size=1000, hidden_size=256, num_layers=4, num_heads=8, intermediate_size=1024,
                             num_kv_blocks=256, block_size=16, max_num_seqs=16, max_num_batched_tokens=64, device="cpu")
    ps = prompts()
    eng.generate(ps, max_new_tokens=20)
    t0 = time.perf_counter()
    outs = eng.generate(ps, max_new_tokens=20)
    dt = time.perf_counter() - t0
    out_tok = sum(len(v) for v in outs.values())
    print(json.dumps({"engine": "reference_as_shipped", "device": "cpu", "dtype": "float32", "hidden": 256, "heads": 8,
                      "kv_blocks": 256, "requests": len(ps), "output_tokens": out_tok, "seconds": dt,
                      "output_tok_per_s": out_tok / dt, "cores": torch.get_num_threads()}), flush=True)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--reference", action="store_true")
    a = ap.parse_args()
    if a.reference:
        run_reference()
    else:
        run_tiny("cpu", OracleOps, torch.float32, 256, 8, 40, "tiny_cpu_oracle")
        if torch.cuda.is_available():
            from vllm_from_scratch_b200 import paged_attention as P
            for graphs in (False, True):
                run_tiny("cuda", P, torch.float32, 256, 8, 40, "tiny_cuda", graphs)
                run_tiny("cuda", P, torch.bfloat16, 1024, 8, 40, "tiny_cuda", graphs)
