"""Pins the CPU oracle (oracle/paged_attention_oracle.py) against the reference:
 * the reference's own known-answer tests, transcribed
   (01_paged_attention/test_paged_attention.py, BLOG_PART1.md:549-663);
 * golden outputs produced by the unmodified reference (tests/golden/make_golden.py).
CPU only; runs in seconds.
"""
import json
import math
import os

import numpy as np
import pytest
import torch

from cases import (DECODE_CASES, PREFILL_CASES, ROPE_CASES, make_decode_inputs, make_prefill_inputs,
                   make_rope_inputs, tensor_sha)
from oracle import paged_attention_oracle as O

GOLD = os.path.join(os.path.dirname(__file__), "golden")
DEC = np.load(os.path.join(GOLD, "decode_cases.npz"))
PRE = np.load(os.path.join(GOLD, "prefill_cases.npz"))
INTS = json.load(open(os.path.join(GOLD, "integer_cases.json")))
ROPE = np.load(os.path.join(GOLD, "rope_cases.npz"))


# ---- integer contracts -----------------------------------------------------
def test_slot_mapping_kat_reference_test():
    # test_paged_attention.py:44-55
    s = O.compute_slot_mapping([3, 7], 6, 4)
    assert (s[0], s[3], s[4], s[5]) == (12, 15, 28, 29)


def test_slot_mapping_kat_blog():
    # BLOG_PART1.md:612-627
    s = O.compute_slot_mapping([3, 7, 1], 10, 4)
    assert (s[0], s[3], s[4], s[8]) == (12, 15, 28, 4)


@pytest.mark.parametrize("c", INTS["slot_mapping"], ids=lambda c: f"bs{c['block_size']}n{c['seq_len']}")
def test_slot_mapping_golden(c):
    assert O.compute_slot_mapping(c["block_table"], c["seq_len"], c["block_size"]) == c["slots"]


def test_pool_allocator_trace():
    tr = INTS["pool_trace"]
    a = O.PoolAllocatorOracle(tr["num_blocks"])
    for ev in tr["trace"]:
        if ev["op"] == "free":
            a.free(ev["ids"])
        elif "error" in ev:
            with pytest.raises(RuntimeError, match="OOM"):
                a.allocate(ev["n"])
            continue
        else:
            assert a.allocate(ev["n"]) == ev["ids"]
        assert a.free_blocks == ev["free_blocks"]
        assert a.ref_counts == ev["ref_counts"]


def test_scheduler_trace_slots():
    tr = INTS["scheduler_trace"]
    for step in tr["steps"]:
        for s in step:
            got = O.compute_slot_mapping_range(s["block_table"], s["positions"][0], len(s["positions"]),
                                               tr["block_size"])
            assert got == s["slots"]


# ---- decode -------------------------------------------------------------------
@pytest.mark.parametrize("case", DECODE_CASES, ids=lambda c: c[0])
def test_decode_oracles_vs_reference_golden(case):
    name, dt, hq, hkv, d, bs, nb, ctx, seed = case
    q, kv, bt = make_decode_inputs(case)
    assert tensor_sha(q, kv, np.asarray(bt, dtype=np.int32)) == str(DEC[f"{name}/in_sha"]), "RNG drift"
    ref = torch.from_numpy(DEC[f"{name}/out"])
    # same-dtype restatements reproduce the reference's numbers (same torch ops, same order)
    if ctx <= 300:
        loop = O.decode_loop(q, kv, bt, ctx, d, hq, hkv, bs).float()
        assert torch.equal(loop, ref)
    vec = O.decode_vectorised(q, kv, bt, ctx, d, hq, hkv, bs).float()
    tol = 1e-6 if dt == "f32" else 2e-2
    assert (vec - ref).abs().max().item() <= tol
    # high-precision oracle: within the north-star tolerances of the reference
    hi = O.decode_vectorised(q, kv, bt, ctx, d, hq, hkv, bs, compute_dtype=torch.float64).float()
    assert (hi - ref).abs().max().item() <= (1e-5 if dt == "f32" else 2e-2)


def test_decode_equals_dense_attention():
    # test_paged_attention.py:67-100, transcribed against the oracle
    torch.manual_seed(42)
    h, d, n, bs = 4, 32, 8, 4
    K, V, Q = torch.randn(n, h, d), torch.randn(n, h, d), torch.randn(h, d)
    sc = torch.einsum('hd,shd->hs', Q, K) / math.sqrt(d)
    std = torch.einsum('hs,shd->hd', torch.softmax(sc, -1), V)
    kv = torch.zeros(2, 4, h, bs, d)
    bt = [2, 0]
    O.write_kv_hbsd(kv, bt, K, V, bs)
    for fn in (O.decode_loop, O.decode_vectorised):
        assert torch.allclose(std, fn(Q, kv, bt, n, d, h, h, bs), atol=1e-5)


def test_placement_invariance():
    # BLOG_PART1.md:629-663
    torch.manual_seed(0)
    h, d, n, bs = 2, 16, 6, 2
    K, V, Q = torch.randn(n, h, d), torch.randn(n, h, d), torch.randn(h, d)
    outs = []
    for ids in ([0, 1, 2], [5, 2, 7]):
        kv = torch.zeros(2, 8, h, bs, d)
        O.write_kv_hbsd(kv, ids, K, V, bs)
        outs.append(O.decode_vectorised(Q, kv, ids, n, d, h, h, bs))
    assert torch.allclose(outs[0], outs[1], atol=1e-5)


def test_decode_bshd_matches_hbsd():
    g = torch.Generator().manual_seed(5)
    nb, bs, h, d = 12, 4, 3, 16
    kc, vc = torch.randn(nb, bs, h, d, generator=g), torch.randn(nb, bs, h, d, generator=g)
    q = torch.randn(2, h, d, generator=g)
    bts = torch.tensor([[3, 1, 7], [2, 9, 0]], dtype=torch.int32)
    lens = torch.tensor([9, 12], dtype=torch.int32)
    got = O.decode_batched_bshd(q, kc, vc, bts, lens)
    for i in range(2):
        n = int(lens[i])
        rows = O.compute_slot_mapping(bts[i].tolist(), n, bs)
        k = kc.view(-1, h, d)[rows]
        v = vc.view(-1, h, d)[rows]
        sc = torch.einsum('hd,shd->hs', q[i], k) / math.sqrt(d)
        ref = torch.einsum('hs,shd->hd', torch.softmax(sc, -1), v)
        assert torch.allclose(got[i], ref, atol=1e-6)


# ---- prefill / writes ---------------------------------------------------------------
@pytest.mark.parametrize("case", PREFILL_CASES, ids=lambda c: c[0])
def test_prefill_oracle_vs_reference_golden(case):
    name, dt, hq, hkv, d, bs, nb, t, seed = case
    q, k, v, kv, bt = make_prefill_inputs(case)
    assert tensor_sha(q, k, v, np.asarray(bt, dtype=np.int32)) == str(PRE[f"{name}/in_sha"]), "RNG drift"
    out = O.prefill(q, k, v, kv, bt, bs)
    assert tensor_sha(kv) == str(PRE[f"{name}/kv_after_sha"])          # bit-exact cache write
    ref = torch.from_numpy(PRE[f"{name}/out"])
    assert torch.equal(out.float(), ref)                                   # same ops, same order
    # slot-indexed scatter writes the same bytes
    kv2 = torch.zeros_like(kv)
    slots = torch.tensor(O.compute_slot_mapping(bt, t, bs))
    O.reshape_and_cache_hbsd(kv2, k, v, slots)
    assert torch.equal(kv2, kv)
    # chunked-prefill oracle with start_pos = 0 equals the reference prefill
    ch = O.chunked_prefill_hbsd(q, kv, bt, 0, bs, compute_dtype=torch.float64).float()
    assert (ch - ref).abs().max().item() <= (1e-5 if dt == "f32" else 2e-2)


def test_chunked_prefill_equals_full_prefill():
    g = torch.Generator().manual_seed(3)
    t, hq, hkv, d, bs = 23, 4, 2, 16, 4
    q, k, v = (torch.randn(t, h, d, generator=g) for h in (hq, hkv, hkv))
    bt = [5, 1, 4, 0, 2, 3]
    kv = torch.zeros(2, 6, hkv, bs, d)
    full = O.prefill(q, k, v, kv, bt, bs)
    parts, start = [], 0
    for n in (7, 1, 10, 5):
        parts.append(O.chunked_prefill_hbsd(q[start:start + n], kv, bt, start, bs))
        start += n
    assert torch.allclose(torch.cat(parts), full, atol=1e-5)


def test_reshape_and_cache_bshd_and_block_ops():
    g = torch.Generator().manual_seed(1)
    nb, bs, h, d, t = 6, 4, 2, 8, 9
    kc, vc = torch.zeros(nb, bs, h, d), torch.zeros(nb, bs, h, d)
    k, v = torch.randn(t, h, d, generator=g), torch.randn(t, h, d, generator=g)
    slots = torch.tensor([20, 21, 22, 23, 4, 5, -1, 6, 7])
    O.reshape_and_cache_bshd(kc, vc, k, v, slots)
    assert torch.equal(kc.view(-1, h, d)[20], k[0]) and torch.equal(vc.view(-1, h, d)[7], v[8])
    assert kc.view(-1, h, d)[0].abs().sum() == 0
    before = kc.clone()
    O.copy_blocks(kc, [(5, 0), (1, 2), (0, 3)], block_dim=0)       # (0,3) must read the OLD block 0
    assert torch.equal(kc[0], before[5]) and torch.equal(kc[2], before[1]) and torch.equal(kc[3], before[0])
    host = torch.zeros(3, bs, h, d)
    O.swap_blocks(kc, host, [(5, 1), (1, 0)], block_dim=0)
    assert torch.equal(host[1], kc[5]) and torch.equal(host[0], kc[1])


# ---- RoPE + QKV split + KV scatter (engine steps 1b-3) ------------------------------------
@pytest.mark.parametrize("case", ROPE_CASES, ids=lambda c: c[0])
def test_rope_oracle_vs_reference_golden(case):
    name, dt, h, d, bs, nb, t, max_len, seed = case
    qkv, positions, slots, k_cache, v_cache = make_rope_inputs(case)
    assert tensor_sha(qkv, positions, slots, k_cache, v_cache) == str(ROPE[f"{name}/in_sha"]), "RNG drift"
    cos, sin = (x.to(qkv.dtype) for x in O.rope_tables(d, max_len))
    q = O.rope_split_and_cache_bshd(qkv, positions, cos, sin, k_cache, v_cache, slots, h, h)
    assert torch.equal(q.float(), torch.from_numpy(ROPE[f"{name}/q"]))              # bit-exact
    assert torch.equal(k_cache.float(), torch.from_numpy(ROPE[f"{name}/k_cache"]))
    assert torch.equal(v_cache.float(), torch.from_numpy(ROPE[f"{name}/v_cache"]))
