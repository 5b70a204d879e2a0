"""Generate golden vectors by running the UNMODIFIED reference in the build container.

Usage (build container only - /root/reference does not exist on the GPU box):

    python tests/golden/make_golden.py

Imports ``/root/reference/01_paged_attention/paged_attention.py`` and
``/root/reference/05_mini_vllm/mini_vllm.py`` read-only through ``sys.path`` (nothing is
copied), runs them on seeded inputs and writes

    tests/golden/decode_cases.npz     reference outputs of paged_attention_decode + sha256 of inputs
    tests/golden/prefill_cases.npz    outputs of paged_attention_prefill + sha256 of inputs / written cache
    tests/golden/integer_cases.json   slot mappings, allocator traces, scheduler traces
    tests/golden/rope_cases.npz       RoPE + QKV split + KV scatter (SelfAttention.forward steps 1b-3)

Inputs are regenerated from seeds by tests/golden/cases.py (shared with the tests) and
pinned by sha256; bf16 outputs are stored upcast to float32 (exact).  The files are small
(< 1 MiB total) and are committed; the GPU parity tests and the CPU oracle tests both read
them, so parity with the real reference is pinned without the reference being present.
"""
import json
import os
import random
import sys

import numpy as np
import torch

REF = os.environ.get("PA_REFERENCE_ROOT", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from cases import (DECODE_CASES, PREFILL_CASES, ROPE_CASES, make_decode_inputs, make_prefill_inputs,  # noqa: E402
                   make_rope_inputs, tensor_sha)


def _import_reference():
    sys.path.insert(0, os.path.join(REF, "01_paged_attention"))
    sys.path.insert(0, os.path.join(REF, "05_mini_vllm"))
    import paged_attention as ref_pa  # noqa
    import mini_vllm as ref_engine  # noqa
    return ref_pa, ref_engine


def _np(t):
    return t.detach().to(torch.float32).cpu().numpy()


def gen_decode(ref_pa):
    out = {}
    for case in DECODE_CASES:
        name, dt, hq, hkv, d, bs, nb, ctx, seed = case
        q, kv, bt = make_decode_inputs(case)
        o = ref_pa.paged_attention_decode(q, kv, bt, ctx, d, hq, hkv, block_size=bs)
        assert o.dtype == q.dtype
        out[f"{name}/out"] = _np(o)
        out[f"{name}/in_sha"] = np.asarray(tensor_sha(q, kv, np.asarray(bt, dtype=np.int32)))
    np.savez_compressed(os.path.join(HERE, "decode_cases.npz"), **out)


def gen_prefill(ref_pa):
    out = {}
    for case in PREFILL_CASES:
        name, dt, hq, hkv, d, bs, nb, t, seed = case
        q, k, v, kv, bt = make_prefill_inputs(case)
        in_sha = tensor_sha(q, k, v, np.asarray(bt, dtype=np.int32))
        o = ref_pa.paged_attention_prefill(q, k, v, kv, bt, bs)
        out[f"{name}/out"] = _np(o)
        out[f"{name}/in_sha"] = np.asarray(in_sha)
        out[f"{name}/kv_after_sha"] = np.asarray(tensor_sha(kv))
    np.savez_compressed(os.path.join(HERE, "prefill_cases.npz"), **out)


def gen_integers(ref_pa, ref_engine):
    js = {}
    # slot mapping: the reference's own KATs + random ones
    slot_cases = [([3, 7], 6, 4), ([3, 7, 1], 10, 4), ([0], 1, 16), ([5, 2, 7], 6, 2)]
    rng = random.Random(0)
    for _ in range(8):
        bs = rng.choice([1, 2, 4, 8, 16, 32])
        n = rng.randint(1, 200)
        nblk = (n + bs - 1) // bs
        slot_cases.append((rng.sample(range(4096), nblk), n, bs))
    js["slot_mapping"] = [
        {"block_table": bt, "seq_len": n, "block_size": bs,
         "slots": ref_pa.compute_slot_mapping(bt, n, bs)} for bt, n, bs in slot_cases]

    # allocator order trace on KVCachePool (tiny tensor; only the lists matter)
    pool = ref_pa.KVCachePool(num_blocks=12, block_size=4, num_kv_heads=1, head_dim=4, num_layers=1)
    trace, held = [], []
    for step in range(40):
        if held and rng.random() < 0.45:
            ids = held.pop(rng.randrange(len(held)))
            pool.free(ids)
            trace.append({"op": "free", "ids": ids, "free_blocks": list(pool.free_blocks),
                          "ref_counts": list(pool.ref_counts)})
        else:
            n = rng.randint(1, 4)
            try:
                ids = pool.allocate(n)
                held.append(ids)
                trace.append({"op": "alloc", "n": n, "ids": ids, "free_blocks": list(pool.free_blocks),
                              "ref_counts": list(pool.ref_counts)})
            except RuntimeError as e:
                trace.append({"op": "alloc", "n": n, "error": str(e)})
    js["pool_trace"] = {"num_blocks": 12, "trace": trace}

    # scheduler trace: block tables + slot mappings produced by the real Scheduler
    random.seed(0)
    alloc = ref_engine.BlockAllocator(num_blocks=64, block_size=4)
    sched = ref_engine.Scheduler(max_num_seqs=4, max_num_batched_tokens=24, block_size=4, allocator=alloc)
    prompts = [[rng.randint(1, 99) for _ in range(rng.randint(3, 40))] for _ in range(7)]
    for i, p in enumerate(prompts):
        sched.add_request(ref_engine.Request(request_id=f"r{i}", prompt_token_ids=p, max_new_tokens=5))
    steps = []
    while sched.has_unfinished and len(steps) < 200:
        out = sched.schedule()
        if not out:
            break
        steps.append([{"rid": s.request.request_id, "positions": s.positions, "slots": s.slot_mapping,
                       "is_prefill": s.is_prefill, "block_table": list(s.request.block_table)} for s in out])
        sched.update(out, [rng.randint(1, 99) for _ in out])
    js["scheduler_trace"] = {"block_size": 4, "prompts": prompts, "steps": steps}
    with open(os.path.join(HERE, "integer_cases.json"), "w") as f:
        json.dump(js, f)


def gen_rope(ref_engine):
    """Steps 1b-3 of the reference SelfAttention.forward (mini_vllm.py:410-427) run with the
    reference's own RotaryEmbedding module and rotate_half on seeded inputs."""
    out = {}
    for case in ROPE_CASES:
        name, dt, h, d, bs, nb, t, max_len, seed = case
        qkv, positions, slots, k_cache, v_cache = make_rope_inputs(case)
        in_sha = tensor_sha(qkv, positions, slots, k_cache, v_cache)
        rope = ref_engine.RotaryEmbedding(d, max_len=max_len).to(qkv.dtype)
        q, k, v = qkv.chunk(3, dim=-1)
        q, k, v = q.view(t, h, d), k.view(t, h, d), v.view(t, h, d)
        cos, sin = rope(positions)
        cos, sin = cos.unsqueeze(1), sin.unsqueeze(1)
        q = q * cos + ref_engine.rotate_half(q) * sin
        k = k * cos + ref_engine.rotate_half(k) * sin
        k_cache.view(-1, h, d)[slots] = k
        v_cache.view(-1, h, d)[slots] = v
        out[f"{name}/q"] = _np(q)
        out[f"{name}/k_cache"] = _np(k_cache)
        out[f"{name}/v_cache"] = _np(v_cache)
        out[f"{name}/in_sha"] = np.asarray(in_sha)
    np.savez_compressed(os.path.join(HERE, "rope_cases.npz"), **out)


if __name__ == "__main__":
    ref_pa, ref_engine = _import_reference()
    if "--only-rope" not in sys.argv:
        gen_decode(ref_pa)
        gen_prefill(ref_pa)
        gen_integers(ref_pa, ref_engine)
    gen_rope(ref_engine)
    for fn in ("decode_cases.npz", "prefill_cases.npz", "integer_cases.json", "rope_cases.npz"):
        print(fn, os.path.getsize(os.path.join(HERE, fn)), "bytes")
