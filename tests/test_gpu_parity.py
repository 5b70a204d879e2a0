"""GPU parity tests proper: the CUDA path (through the Python host -> C ABI of libpa_b200.so)
against (a) the reference's own outputs stored in tests/golden/, (b) the reference's unit tests
replayed on CUDA tensors (01_paged_attention/test_paged_attention.py), (c) the CPU oracle on
the same seeded inputs, and (d) size-independent properties at BASELINE.json's full sizes.

Tolerances (BASELINE.json north_star): attention max-abs <= 2e-2 in bf16, <= 1e-5 with an
fp32 KV cache; slot mapping / KV writes / block copies / swaps bit-exact.
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

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(__file__), "golden")
TOL = {"f32": 1e-5, "bf16": 2e-2, torch.float32: 1e-5, torch.bfloat16: 2e-2}


@pytest.fixture(scope="module")
def P():
    from vllm_from_scratch_b200 import paged_attention as mod
    from vllm_from_scratch_b200 import _lib
    _lib.load()
    return mod


@pytest.fixture(scope="module")
def lib():
    from vllm_from_scratch_b200 import _lib
    return _lib.load()


@pytest.fixture(params=[1, 2], ids=["stream", "plain"])
def force_generic(lib, request):
    """1 = no TMA kernel (the cp.async streaming kernel where rows are 16-byte aligned), 2 = the plain generic kernel."""
    old = lib.pa_decode_force_path(request.param)
    yield
    lib.pa_decode_force_path(old)


DEV = "cuda"


# =========================================================================================
# (b) the reference's unit tests, replayed on CUDA (test_paged_attention.py)
# =========================================================================================
@pytest.fixture()
def small_pool(P):
    # test_paged_attention.py:14-18
    return P.KVCachePool(num_blocks=16, block_size=16, num_kv_heads=4, head_dim=32, num_layers=2,
                         dtype=torch.float32, device=DEV)


def test_ref_alloc_free_basic(small_pool):            # :22-28
    blocks = small_pool.allocate(4)
    assert len(blocks) == 4 and len(small_pool.free_blocks) == 12
    small_pool.free(blocks)
    assert len(small_pool.free_blocks) == 16


def test_ref_oom(small_pool):                          # :30-32
    with pytest.raises(RuntimeError, match="OOM"):
        small_pool.allocate(100)


def test_ref_write_read_kv(small_pool):                # :34-40
    k, v = torch.randn(4, 32, device=DEV), torch.randn(4, 32, device=DEV)
    small_pool.write_kv(0, 0, 5, k, v)
    assert torch.allclose(small_pool.kv_cache[0, 0, 0, :, 5, :], k)
    assert torch.allclose(small_pool.kv_cache[1, 0, 0, :, 5, :], v)
    assert torch.equal(small_pool.kv_cache[0, 0, 0, :, 5, :], k)      # stronger: bit-exact
    assert small_pool.kv_cache.abs().sum().item() == pytest.approx((k.abs().sum() + v.abs().sum()).item(), rel=1e-5)


def test_ref_slot_mapping(P):                          # :44-55
    slots = P.compute_slot_mapping([3, 7], 6, 4)
    assert (slots[0], slots[3], slots[4], slots[5]) == (12, 15, 28, 29)
    dev = P.compute_slot_mapping_device(torch.tensor([3, 7], dtype=torch.int32, device=DEV), 0, 6, 4)
    assert dev.tolist() == slots


def test_ref_decode_shape(P):                          # :59-65
    kv = torch.randn(2, 8, 4, 16, 32, device=DEV)
    q = torch.randn(8, 32, device=DEV)
    out = P.paged_attention_decode(q, kv, [0, 1], 20, 32, 8, 4, 16)
    assert out.shape == (8, 32)


def test_ref_decode_equals_dense_attention(P):         # :67-100
    torch.manual_seed(42)
    h, d, n, bs = 4, 32, 8, 4
    K, V, Q = torch.randn(n, h, d), torch.randn(n, h, d), torch.randn(h, d)
    sc = torch.einsum('hd,shd->hs', Q, K) / math.sqrt(d)
    std = torch.einsum('hs,shd->hd', torch.softmax(sc, -1), V)
    kv = torch.zeros(2, 4, h, bs, d)
    bt = [2, 0]
    for i in range(n):
        kv[0, bt[i // bs], :, i % bs] = K[i]
        kv[1, bt[i // bs], :, i % bs] = V[i]
    out = P.paged_attention_decode(Q.to(DEV), kv.to(DEV), bt, n, d, h, h, bs)
    assert torch.allclose(std, out.cpu(), atol=1e-5)


def test_ref_gqa_shape(P):                             # :102-110
    kv = torch.randn(2, 2, 2, 16, 16, device=DEV)
    q = torch.randn(8, 16, device=DEV)
    out = P.paged_attention_decode(q, kv, [0], 4, 16, 8, 2, 16)
    assert out.shape == (8, 16)
    ref = O.decode_vectorised(q.cpu(), kv.cpu(), [0], 4, 16, 8, 2, 16, compute_dtype=torch.float64)
    assert (out.cpu().double() - ref).abs().max().item() <= 1e-5       # values too, not only shape


def test_ref_prefill_writes_cache_and_shape(P):        # :114-137
    t, h, d, bs = 4, 4, 16, 4
    q, k, v = (torch.randn(t, h, d, device=DEV) for _ in range(3))
    kv = torch.zeros(2, 2, h, bs, d, device=DEV)
    out = P.paged_attention_prefill(q, k, v, kv, [0], bs)
    assert torch.allclose(kv[0, 0, :, 0], k[0]) and torch.allclose(kv[0, 0, :, 3], k[3])
    assert out.shape == (t, h, d)


def test_blog_slot_mapping_and_placement_invariance(P):    # BLOG_PART1.md:612-663
    s = P.compute_slot_mapping([3, 7, 1], 10, 4)
    assert (s[0], s[3], s[4], s[8]) == (12, 15, 28, 4)
    torch.manual_seed(0)
    h, d, n, bs = 2, 16, 6, 2
    K, V, Q = torch.randn(n, h, d), torch.randn(n, h, d), torch.randn(h, d)
    outs = []
    for ids in ([0, 1, 2], [5, 2, 7]):
        kv = torch.zeros(2, 8, h, bs, d, device=DEV)
        P.paged_attention_prefill(Q.new_zeros(n, h, d).to(DEV), K.to(DEV), V.to(DEV), kv, ids, bs)
        outs.append(P.paged_attention_decode(Q.to(DEV), kv, ids, n, d, h, h, bs))
    assert torch.equal(outs[0], outs[1])        # same logical order -> identical arithmetic


# =========================================================================================
# (a) reference outputs (golden) - decode
# =========================================================================================
DEC = np.load(os.path.join(GOLD, "decode_cases.npz"))
PRE = np.load(os.path.join(GOLD, "prefill_cases.npz"))


def _run_decode_case(P, case):
    name, dt, hq, hkv, d, bs, nb, ctx, seed = case
    q, kv, bt = make_decode_inputs(case)
    assert tensor_sha(q, kv, np.asarray(bt, dtype=np.int32)) == str(DEC[f"{name}/in_sha"]), "RNG drift"
    out = P.paged_attention_decode(q.to(DEV), kv.to(DEV), bt, ctx, d, hq, hkv, bs)
    assert out.dtype == q.dtype and out.shape == (hq, d)
    ref = torch.from_numpy(DEC[f"{name}/out"])                      # the reference's own output
    err = (out.float().cpu() - ref).abs().max().item()
    assert err <= TOL[dt], f"{name}: max-abs {err} vs reference"
    hi = O.decode_vectorised(q, kv, bt, ctx, d, hq, hkv, bs, compute_dtype=torch.float64)
    err64 = (out.double().cpu() - hi).abs().max().item()
    assert err64 <= TOL[dt], f"{name}: max-abs {err64} vs fp64 oracle"


@pytest.mark.parametrize("case", DECODE_CASES, ids=lambda c: c[0])
def test_decode_golden(P, case):
    _run_decode_case(P, case)


@pytest.mark.parametrize("case", DECODE_CASES, ids=lambda c: c[0])
def test_decode_golden_generic_kernel(P, force_generic, case):
    _run_decode_case(P, case)


@pytest.mark.parametrize("case", PREFILL_CASES, ids=lambda c: c[0])
def test_prefill_golden(P, case):
    name, dt, hq, hkv, d, bs, nb, t, seed = case
    q, k, v, kv, bt = make_prefill_inputs(case)
    assert tensor_sha(q, k, v, np.asarray(bt, dtype=np.int32)) == str(PRE[f"{name}/in_sha"]), "RNG drift"
    kv_d = kv.to(DEV)
    out = P.paged_attention_prefill(q.to(DEV), k.to(DEV), v.to(DEV), kv_d, bt, bs)
    assert tensor_sha(kv_d) == str(PRE[f"{name}/kv_after_sha"])         # cache write bit-exact vs reference
    ref = torch.from_numpy(PRE[f"{name}/out"])
    err = (out.float().cpu() - ref).abs().max().item()
    assert err <= TOL[dt], f"{name}: max-abs {err}"


# =========================================================================================
# (c) oracle on seeded inputs: batched / ragged / both layouts / both kernels
# =========================================================================================
def _make_batch(b, hq, hkv, d, bs, lens, dtype, seed, nb_extra=3):
    g = torch.Generator().manual_seed(seed)
    nblk = [(n + bs - 1) // bs for n in lens]
    nb = sum(nblk) + nb_extra
    kv = torch.randn(2, nb, hkv, bs, d, generator=g).to(dtype)
    q = torch.randn(b, hq, d, generator=g).to(dtype)
    perm = torch.randperm(nb, generator=g).tolist()
    maxb = max(max(nblk), 1)
    bts = torch.full((b, maxb + 2), -12345, dtype=torch.int32)      # poison: must never be dereferenced
    at = 0
    for i, n in enumerate(nblk):
        bts[i, :n] = torch.tensor(perm[at:at + n], dtype=torch.int32)
        at += n
    bts_safe = bts.clamp(min=0)
    return q, kv, bts, bts_safe, torch.tensor(lens, dtype=torch.int32)


RAGGED = [
    # (hq, hkv, d, bs, dtype, lens)
    (32, 8, 128, 16, torch.bfloat16, [1, 15, 16, 17, 0, 300, 129, 2048, 33]),
    (32, 8, 128, 16, torch.float32, [1, 15, 16, 17, 0, 300, 129, 700]),
    (8, 8, 64, 16, torch.bfloat16, [5, 64, 0, 0, 1000, 31]),
    (16, 1, 128, 16, torch.bfloat16, [257, 3, 48]),
    (12, 4, 32, 4, torch.float32, [1, 2, 3, 4, 5, 9, 0, 41]),
    (6, 2, 80, 8, torch.bfloat16, [7, 8, 9, 100]),                   # D not a power of two -> generic
    (40, 4, 64, 16, torch.bfloat16, [100, 17]),                       # g = 10 -> two n-tiles
    (24, 1, 32, 16, torch.float32, [50, 16]),                         # g = 24 -> three q tiles (generic)
    (4, 4, 16, 2, torch.float32, [0]),                                # nothing but an empty sequence
]


@pytest.mark.parametrize("cfg", RAGGED, ids=lambda c: f"hq{c[0]}kv{c[1]}d{c[2]}bs{c[3]}{'bf16' if c[4] == torch.bfloat16 else 'f32'}")
@pytest.mark.parametrize("path", ["auto", "generic", "plain"])
def test_decode_batched_ragged_vs_oracle(P, lib, cfg, path):
    hq, hkv, d, bs, dtype, lens = cfg
    q, kv, bts, bts_safe, sl = _make_batch(len(lens), hq, hkv, d, bs, lens, dtype, seed=len(lens) * 7 + d)
    old = lib.pa_decode_force_path({"auto": 0, "generic": 1, "plain": 2}[path])
    try:
        out = P.paged_attention_decode_batched(q.to(DEV), kv.to(DEV), bts.to(DEV), sl.to(DEV))
    finally:
        lib.pa_decode_force_path(old)
    ref = O.decode_batched(q, kv, bts_safe, sl, bs, compute_dtype=torch.float64)
    err = (out.double().cpu() - ref).abs().max().item()
    assert err <= TOL[dtype], f"max-abs {err}"
    for i, n in enumerate(lens):
        if n == 0:
            assert out[i].abs().sum().item() == 0          # defined extension: empty context -> zeros
    # the reference's same-dtype arithmetic agrees to the same tolerance
    same = torch.stack([O.decode_vectorised(q[i], kv, bts_safe[i].tolist(), n, d, hq, hkv, bs).float()
                        for i, n in enumerate(lens)])
    assert (out.float().cpu() - same).abs().max().item() <= TOL[dtype]


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("shape", [(8, 32, 16), (8, 128, 16), (4, 64, 8)])
def test_decode_bshd_engine_layout(P, dtype, shape):
    # mini_vllm.py:544-552 layout [NB, BS, H, D], no GQA
    h, d, bs = shape
    g = torch.Generator().manual_seed(11)
    lens = [1, 17, 40, 0, 64]
    nblk = [(n + bs - 1) // bs for n in lens]
    nb = sum(nblk) + 2
    kc = torch.randn(nb, bs, h, d, generator=g).to(dtype)
    vc = torch.randn(nb, bs, h, d, generator=g).to(dtype)
    q = torch.randn(len(lens), h, d, generator=g).to(dtype)
    perm = torch.randperm(nb, generator=g).tolist()
    bts = torch.zeros(len(lens), max(nblk), dtype=torch.int32)
    at = 0
    for i, n in enumerate(nblk):
        bts[i, :n] = torch.tensor(perm[at:at + n], dtype=torch.int32)
        at += n
    sl = torch.tensor(lens, dtype=torch.int32)
    out = P.paged_attention_decode_batched(q.to(DEV), (kc.to(DEV), vc.to(DEV)), bts.to(DEV), sl.to(DEV))
    ref = O.decode_batched_bshd(q, kc, vc, bts, sl, compute_dtype=torch.float64)
    assert (out.double().cpu() - ref).abs().max().item() <= TOL[dtype]


def test_decode_pool_layer_view_and_strided_query(P):
    # per-layer slice of the 6-D pool (non-contiguous plane stride) and a strided query view
    pool = P.KVCachePool(num_blocks=40, block_size=16, num_kv_heads=8, head_dim=128, num_layers=3,
                         dtype=torch.bfloat16, device=DEV)
    g = torch.Generator().manual_seed(2)
    pool.kv_cache.copy_(torch.randn(pool.kv_cache.shape, generator=g).to(torch.bfloat16))
    qkv = torch.randn(3, 48, 128, generator=g).to(torch.bfloat16).to(DEV)      # fused [B, Hq + 2 Hkv, D]
    q = qkv[:, :32]
    blocks = pool.allocate(30)
    bts = torch.tensor([blocks[0:10], blocks[10:20], blocks[20:30]], dtype=torch.int32)
    sl = torch.tensor([160, 145, 7], dtype=torch.int32)
    for layer in range(3):
        out = P.paged_attention_decode_batched(q, pool.layer(layer), bts.to(DEV), sl.to(DEV))
        ref = O.decode_batched(q.cpu(), pool.layer(layer).cpu(), bts, sl, 16, compute_dtype=torch.float64)
        assert (out.double().cpu() - ref).abs().max().item() <= 2e-2


# =========================================================================================
# (d) BASELINE.json full sizes
# =========================================================================================
def _full_case(b, ctx_lo, ctx_hi, seed, nb_spare=64, hq=32, hkv=8, d=128, bs=16):
    g = torch.Generator(device=DEV).manual_seed(seed)
    gc = torch.Generator().manual_seed(seed)
    if ctx_lo == ctx_hi:
        lens = torch.full((b,), ctx_hi, dtype=torch.int32)
    else:
        lens = torch.randint(ctx_lo, ctx_hi + 1, (b,), generator=gc, dtype=torch.int32)
    nblk = (lens + bs - 1) // bs
    nb = int(nblk.sum()) + nb_spare
    kv = torch.randn(2, nb, hkv, bs, d, generator=g, device=DEV, dtype=torch.float32).to(torch.bfloat16)
    q = torch.randn(b, hq, d, generator=g, device=DEV, dtype=torch.float32).to(torch.bfloat16)
    perm = torch.randperm(nb, generator=gc).to(torch.int32)
    bts = torch.zeros(b, int(nblk.max()), dtype=torch.int32)
    at = 0
    for i in range(b):
        n = int(nblk[i])
        bts[i, :n] = perm[at:at + n]
        at += n
    return q, kv, bts, lens


@pytest.mark.parametrize("ragged", [False, True], ids=["equal", "ragged"])
def test_c2_llama_decode_64x2k(P, ragged):
    # BASELINE.json configs[1]: 64 seqs x 2k, 32/8 heads, D 128, bf16, BS 16
    q, kv, bts, lens = _full_case(64, 1024 if ragged else 2048, 2048, seed=0)
    out = P.paged_attention_decode_batched(q, kv, bts.to(DEV), lens.to(DEV))
    kv_c, q_c = kv.cpu(), q.cpu()
    ref = O.decode_batched(q_c, kv_c, bts, lens, 16, compute_dtype=torch.float32)
    err = (out.float().cpu() - ref).abs().max().item()
    assert err <= 2e-2, f"max-abs {err}"
    # two sequences against the reference's row-by-row algorithm in its own bf16 arithmetic
    for i in (0, 63):
        n = min(int(lens[i]), 400)                              # keep the Python loop to seconds
        loop = O.decode_loop(q_c[i], kv_c, bts[i].tolist(), n, 128, 32, 8, 16)
        lens1 = torch.tensor([n], dtype=torch.int32, device=DEV)
        got = P.paged_attention_decode_batched(q[i:i + 1], kv, bts[i:i + 1].to(DEV), lens1)[0]
        assert (got.float().cpu() - loop.float()).abs().max().item() <= 2e-2
    # determinism: same launch twice -> same bits
    out2 = P.paged_attention_decode_batched(q, kv, bts.to(DEV), lens.to(DEV))
    assert torch.equal(out, out2)


def test_c3_long_context_8x32k_properties(P):
    # BASELINE.json configs[2]: 8 seqs x 32k.  Oracle on 2 sequences; properties on all.
    q, kv, bts, lens = _full_case(8, 32768, 32768, seed=1, nb_spare=16)
    out = P.paged_attention_decode_batched(q, kv, bts.to(DEV), lens.to(DEV))
    kv_c = kv.cpu()
    for i in (0, 7):
        ref = O.decode_vectorised(q[i].cpu(), kv_c, bts[i].tolist(), 32768, 128, 32, 8, 16,
                                  compute_dtype=torch.float32)
        assert (out[i].float().cpu() - ref).abs().max().item() <= 2e-2
    # property 1: physical placement invariance - move every block, outputs keep their bits
    nb = kv.shape[1]
    perm = torch.randperm(nb, generator=torch.Generator().manual_seed(5))
    inv = torch.empty_like(perm)
    inv[perm] = torch.arange(nb)
    kv2 = kv[:, perm.to(DEV)]                      # new block j holds old block perm[j]
    bts2 = inv[bts.long()].to(torch.int32)          # old id -> new id
    out_moved = P.paged_attention_decode_batched(q, kv2, bts2.to(DEV), lens.to(DEV))
    assert torch.equal(out, out_moved)
    # property 2: a convex combination - every output lies inside the per-dim range of V
    vmin = kv[1].float().amin(dim=(0, 2))          # [Hkv, D]
    vmax = kv[1].float().amax(dim=(0, 2))
    o = out.float().view(8, 8, 4, 128)
    assert bool(((o >= vmin[None, :, None] - 1e-2) & (o <= vmax[None, :, None] + 1e-2)).all())
    # property 3: batch independence - a sub-batch gives the same bits for its sequences?  (the
    # split points move, so only tolerance-equality is promised)
    sub = P.paged_attention_decode_batched(q[2:5], kv, bts[2:5].to(DEV), lens[2:5].to(DEV))
    assert (sub.float() - out[2:5].float()).abs().max().item() <= 2e-2


def test_c5_shard_64x8k_properties(P):
    # BASELINE.json configs[4], one GPU's shard of 512 x 8k over 8 GPUs: 64 seqs x 8192, ragged tail.
    q, kv, bts, lens = _full_case(64, 8192 - 300, 8192, seed=11, nb_spare=16)
    lens_d, bts_d = lens.to(DEV), bts.to(DEV)
    out = P.paged_attention_decode_batched(q, kv, bts_d, lens_d)
    kv_c = kv.cpu()
    for i in (0, 31, 63):                                      # oracle on sampled sequences
        ref = O.decode_vectorised(q[i].cpu(), kv_c, bts[i].tolist(), int(lens[i]), 128, 32, 8, 16,
                                  compute_dtype=torch.float32)
        assert (out[i].float().cpu() - ref).abs().max().item() <= 2e-2
    assert torch.equal(out, P.paged_attention_decode_batched(q, kv, bts_d, lens_d))       # reproducible bits
    # linearity in V: attention weights do not depend on V, so out(V) + out(V') == out(V + V') up to bf16 rounding
    kv_b = kv.clone()
    kv_b[1] = torch.randn(kv[1].shape, generator=torch.Generator(device=DEV).manual_seed(12), device=DEV).to(torch.bfloat16)
    kv_sum = kv.clone()
    kv_sum[1] = (kv[1].float() + kv_b[1].float()).to(torch.bfloat16)
    out_b = P.paged_attention_decode_batched(q, kv_b, bts_d, lens_d)
    out_sum = P.paged_attention_decode_batched(q, kv_sum, bts_d, lens_d)
    assert (out_sum.float() - (out.float() + out_b.float())).abs().max().item() <= 3e-2
    # a sequence's result does not depend on its neighbours in the batch (tolerance: split points move)
    one = P.paged_attention_decode_batched(q[40:41], kv, bts_d[40:41], lens_d[40:41])
    assert (one.float() - out[40:41].float()).abs().max().item() <= 2e-2


@pytest.mark.parametrize("hkv", [1, 2, 4])
def test_head_shard_rank_shapes_vs_oracle(P, hkv):
    # per-rank shapes of KV-head-sharded 32k decode (Hkv / P heads): few, very long segments - the
    # split-K plan caps the partials per segment (kMaxPartials), every head must still match the oracle
    hq = 4 * hkv
    q, kv, bts, lens = _full_case(3, 32768 - 77, 32768, seed=20 + hkv, nb_spare=8, hq=hq, hkv=hkv)
    out = P.paged_attention_decode_batched(q, kv, bts.to(DEV), lens.to(DEV))
    kv_c = kv.cpu()
    for i in range(3):
        ref = O.decode_vectorised(q[i].cpu(), kv_c, bts[i].tolist(), int(lens[i]), 128, hq, hkv, 16,
                                  compute_dtype=torch.float32)
        assert (out[i].float().cpu() - ref).abs().max().item() <= 2e-2
    assert torch.equal(out, P.paged_attention_decode_batched(q, kv, bts.to(DEV), lens.to(DEV)))


def test_constant_v_property(P):
    # softmax weights sum to one: with V == c everywhere the output is c (any split-K plan)
    q, kv, bts, lens = _full_case(16, 100, 5000, seed=3)
    kv[1].fill_(0.5)
    out = P.paged_attention_decode_batched(q, kv, bts.to(DEV), lens.to(DEV))
    assert (out.float() - 0.5).abs().max().item() <= 4e-3


def test_garbage_past_seq_len_is_ignored(P):
    # slots past seq_len in the last block may hold NaN/Inf (recycled blocks): must not leak
    q, kv, bts, lens = _full_case(4, 1, 200, seed=4, nb_spare=4)
    ref = P.paged_attention_decode_batched(q, kv, bts.to(DEV), lens.to(DEV))
    kv2 = kv.clone()
    for i in range(4):
        n = int(lens[i])
        if n % 16:
            kv2[:, int(bts[i, n // 16]), :, n % 16:, :] = float("nan")
    got = P.paged_attention_decode_batched(q, kv2, bts.to(DEV), lens.to(DEV))
    assert torch.equal(ref, got)
    old = os.environ.get("X")
    from vllm_from_scratch_b200 import _lib
    for path in (1, 2):
        prev = _lib.load().pa_decode_force_path(path)
        try:
            got_g = P.paged_attention_decode_batched(q, kv2, bts.to(DEV), lens.to(DEV))
        finally:
            _lib.load().pa_decode_force_path(prev)
        assert not torch.isnan(got_g).any()
        assert (got_g.float() - ref.float()).abs().max().item() <= 2e-2


# =========================================================================================
# KV slot write, slot mapping, block copy, swap - bit-exact
# =========================================================================================
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("shape", [(8, 128, 16, 1000), (4, 32, 16, 37), (2, 24, 4, 9), (3, 7, 2, 5)])
def test_reshape_and_cache_both_layouts_bit_exact(P, dtype, shape):
    hkv, d, bs, t = shape
    g = torch.Generator().manual_seed(t)
    nb = (t + bs - 1) // bs + 5
    k = torch.randn(t, hkv, d, generator=g).to(dtype)
    v = torch.randn(t, hkv, d, generator=g).to(dtype)
    slots = torch.randperm(nb * bs, generator=g)[:t]
    slots[t // 2] = -1                                         # padding token: skipped
    # 'hbsd'
    kv_ref = torch.randn(2, nb, hkv, bs, d, generator=g).to(dtype)
    kv_dev = kv_ref.to(DEV)
    O.reshape_and_cache_hbsd(kv_ref, k, v, slots)
    P.reshape_and_cache(k.to(DEV), v.to(DEV), kv_dev, slots.to(DEV))
    assert torch.equal(kv_dev.cpu().view(torch.uint8), kv_ref.view(torch.uint8))
    # 'bshd'
    kc, vc = torch.randn(nb, bs, hkv, d, generator=g).to(dtype), torch.randn(nb, bs, hkv, d, generator=g).to(dtype)
    kc_d, vc_d = kc.to(DEV), vc.to(DEV)
    O.reshape_and_cache_bshd(kc, vc, k, v, slots)
    P.reshape_and_cache(k.to(DEV), v.to(DEV), (kc_d, vc_d), slots.to(DEV).to(torch.int32))
    assert torch.equal(kc_d.cpu().view(torch.uint8), kc.view(torch.uint8))
    assert torch.equal(vc_d.cpu().view(torch.uint8), vc.view(torch.uint8))


def test_reshape_and_cache_strided_source(P):
    # K / V as views of a fused qkv projection output [T, (Hq + 2 Hkv), D] (mini_vllm.py:408-414)
    t, hq, hkv, d, bs = 50, 8, 2, 64, 16
    g = torch.Generator().manual_seed(0)
    qkv = torch.randn(t, hq + 2 * hkv, d, generator=g).to(torch.bfloat16)
    k, v = qkv[:, hq:hq + hkv], qkv[:, hq + hkv:]
    kv_ref = torch.zeros(2, 6, hkv, bs, d, dtype=torch.bfloat16)
    slots = torch.tensor(O.compute_slot_mapping([4, 1, 5, 0], t, bs))
    O.reshape_and_cache_hbsd(kv_ref, k, v, slots)
    qkv_d = qkv.to(DEV)
    kv_dev = torch.zeros_like(kv_ref, device=DEV)
    P.reshape_and_cache(qkv_d[:, hq:hq + hkv], qkv_d[:, hq + hkv:], kv_dev, slots.to(DEV))
    assert torch.equal(kv_dev.cpu(), kv_ref)


def test_slot_mapping_device_golden(P):
    ints = json.load(open(os.path.join(GOLD, "integer_cases.json")))
    for c in ints["slot_mapping"]:
        bt = torch.tensor(c["block_table"], dtype=torch.int32, device=DEV)
        got = P.compute_slot_mapping_device(bt, 0, c["seq_len"], c["block_size"]).tolist()
        assert got == c["slots"] == P.compute_slot_mapping(c["block_table"], c["seq_len"], c["block_size"])
    tr = ints["scheduler_trace"]
    for step in tr["steps"]:
        for s in step:
            bt = torch.tensor(s["block_table"], dtype=torch.int32, device=DEV)
            got = P.compute_slot_mapping_device(bt, s["positions"][0], len(s["positions"]), tr["block_size"])
            assert got.tolist() == s["slots"]


def test_pool_allocator_trace(P):
    tr = json.load(open(os.path.join(GOLD, "integer_cases.json")))["pool_trace"]
    pool = P.KVCachePool(tr["num_blocks"], 4, 1, 8, 1, device=DEV)
    for ev in tr["trace"]:
        if ev["op"] == "free":
            pool.free(ev["ids"])
        elif "error" in ev:
            with pytest.raises(RuntimeError, match="OOM"):
                pool.allocate(ev["n"])
            continue
        else:
            assert pool.allocate(ev["n"]) == ev["ids"]
        assert pool.free_blocks == ev["free_blocks"] and pool.ref_counts == ev["ref_counts"]


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_copy_blocks_bit_exact(P, dtype):
    g = torch.Generator().manual_seed(9)
    pool = torch.randn(2, 3, 20, 4, 16, 32, generator=g).to(dtype)         # [2, L, NB, Hkv, BS, D]
    pairs = [(5, 0), (1, 2), (7, 3), (7, 19)]
    ref = pool.clone()
    O.copy_blocks(ref, pairs, block_dim=2)
    dev = pool.to(DEV)
    P.copy_blocks(dev, pairs)
    assert torch.equal(dev.cpu().view(torch.uint8), ref.view(torch.uint8))
    # one layer view + engine layout pair + device-resident pairs tensor
    lay = pool.to(DEV)
    P.copy_blocks(lay[:, 1], torch.tensor(pairs, device=DEV))
    ref2 = pool.clone()
    O.copy_blocks(ref2[:, 1], pairs, block_dim=1)
    assert torch.equal(lay.cpu().view(torch.uint8), ref2.view(torch.uint8))
    kc, vc = torch.randn(2, 9, 4, 2, 8, generator=g).to(dtype), torch.randn(2, 9, 4, 2, 8, generator=g).to(dtype)
    kd, vd = kc.to(DEV), vc.to(DEV)
    P.copy_blocks([kd, vd], [(0, 8), (3, 4)])
    O.copy_blocks(kc, [(0, 8), (3, 4)], 1)
    O.copy_blocks(vc, [(0, 8), (3, 4)], 1)
    assert torch.equal(kd.cpu(), kc) and torch.equal(vd.cpu(), vc)
    with pytest.raises(ValueError):
        P.copy_blocks(dev, [(1, 2), (2, 3)])


def test_swap_out_in_round_trip(P):
    pool = P.KVCachePool(num_blocks=32, block_size=16, num_kv_heads=8, head_dim=128, num_layers=4,
                         dtype=torch.bfloat16, device=DEV)
    g = torch.Generator().manual_seed(4)
    pool.kv_cache.copy_(torch.randn(pool.kv_cache.shape, generator=g).to(torch.bfloat16))
    before = pool.kv_cache.cpu()
    host = P.HostKVPool(pool.kv_cache, num_blocks=8)
    dev_blocks = [3, 4, 5, 17, 30]                       # a run of 3 + two singles
    host_blocks = host.allocate(len(dev_blocks))
    ev = P.swap_out(pool.kv_cache, host.kv_cache, list(zip(dev_blocks, host_blocks)))
    ev.synchronize()
    ref_host = torch.zeros_like(host.kv_cache)
    O.swap_blocks(before, ref_host, list(zip(dev_blocks, host_blocks)), block_dim=2)
    assert torch.equal(host.kv_cache[:, :, host_blocks].view(torch.uint8), ref_host[:, :, host_blocks].view(torch.uint8))
    # erase on the device, bring back into DIFFERENT blocks, attention sees the same values
    pool.kv_cache[:, :, dev_blocks] = 0
    new_blocks = [9, 10, 11, 12, 2]
    ev = P.swap_in(host.kv_cache, pool.kv_cache, list(zip(host_blocks, new_blocks)))
    torch.cuda.current_stream().wait_event(ev)
    after = pool.kv_cache.cpu()
    assert torch.equal(after[:, :, new_blocks].view(torch.uint8), before[:, :, dev_blocks].view(torch.uint8))
    untouched = [b for b in range(32) if b not in dev_blocks + new_blocks]
    assert torch.equal(after[:, :, untouched], before[:, :, untouched])


@pytest.mark.parametrize("layout", ["pool", "layer", "engine"])
@pytest.mark.parametrize("staging_bytes", [None, 96 << 10], ids=["one-chunk", "chunked"])
def test_swap_scattered_blocks_staged_path(P, layout, staging_bytes, monkeypatch):
    # many scattered device blocks <-> consecutive host blocks: the staged path (gather kernel ->
    # dense staging -> one pinned copy per (plane, host run)); a tiny staging cap forces chunking
    if staging_bytes is not None:
        monkeypatch.setattr(P, "_SWAP_STAGING_BYTES", staging_bytes)
    g = torch.Generator().manual_seed(12)
    nb, nh = 96, 48
    if layout == "pool":
        dev = torch.randn(2, 3, nb, 2, 16, 64, generator=g).to(torch.bfloat16).to(DEV)   # [2, L, NB, Hkv, BS, D]
    elif layout == "layer":
        dev = torch.randn(2, 3, nb, 2, 16, 64, generator=g).to(torch.bfloat16).to(DEV)[:, 1]
    else:
        dev = torch.randn(3, nb, 16, 2, 64, generator=g).to(torch.float32).to(DEV)      # [L, NB, BS, H, D]
    bd = dev.dim() - 4
    before = dev.cpu().clone()
    host = P.HostKVPool(dev, num_blocks=nh)
    host.kv_cache.zero_()
    dev_blocks = torch.randperm(nb, generator=g)[:40].tolist()
    host_blocks = host.allocate(30) + [40, 45, 41, 47, 33, 32, 31, 46, 44, 42]       # a long run + scattered tail
    P.swap_out(dev, host.kv_cache, list(zip(dev_blocks, host_blocks))).synchronize()
    ref_host = torch.zeros_like(host.kv_cache)
    O.swap_blocks(before, ref_host, list(zip(dev_blocks, host_blocks)), block_dim=bd)
    assert torch.equal(host.kv_cache.view(torch.uint8), ref_host.view(torch.uint8))
    # back into other scattered blocks
    dev.index_fill_(bd, torch.tensor(dev_blocks, device=DEV), 0)
    new_blocks = torch.randperm(nb, generator=g)[:40].tolist()
    torch.cuda.current_stream().wait_event(P.swap_in(host.kv_cache, dev, list(zip(host_blocks, new_blocks))))
    expect = before.clone()
    expect.index_fill_(bd, torch.tensor(dev_blocks), 0)
    O.swap_blocks(ref_host, expect, list(zip(host_blocks, new_blocks)), block_dim=bd)
    assert torch.equal(dev.cpu().view(torch.uint8), expect.view(torch.uint8))


def test_swap_then_decode_matches(P):
    # decode -> swap the sequence out -> swap it into other blocks -> decode: identical bits
    q, kv, bts, lens = _full_case(2, 500, 900, seed=8, nb_spare=80)
    out0 = P.paged_attention_decode_batched(q, kv, bts.to(DEV), lens.to(DEV))
    host = P.HostKVPool(kv, num_blocks=64)
    n0 = (int(lens[0]) + 15) // 16
    src = bts[0, :n0].tolist()
    hb = host.allocate(n0)
    P.swap_out(kv, host.kv_cache, list(zip(src, hb))).synchronize()
    kv[:, src] = float("nan")
    used = set(bts.flatten().tolist())
    free = [b for b in range(kv.shape[1]) if b not in used][:n0]
    torch.cuda.current_stream().wait_event(P.swap_in(host.kv_cache, kv, list(zip(hb, free))))
    bts2 = bts.clone()
    bts2[0, :n0] = torch.tensor(free, dtype=torch.int32)
    out1 = P.paged_attention_decode_batched(q, kv, bts2.to(DEV), lens.to(DEV))
    assert torch.equal(out0, out1)


# =========================================================================================
# chunked prefill (history through the block table) and error behaviour
# =========================================================================================
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_chunked_prefill_vs_oracle(P, dtype):
    g = torch.Generator().manual_seed(3)
    t, hq, hkv, d, bs = 150, 8, 2, 64, 16
    q, k, v = (torch.randn(t, h, d, generator=g).to(dtype) for h in (hq, hkv, hkv))
    bt = torch.randperm(12, generator=g)[:10].tolist()
    kv_ref = torch.zeros(2, 12, hkv, bs, d, dtype=dtype)
    O.write_kv_hbsd(kv_ref, bt, k, v, bs)
    kv_dev = torch.zeros(2, 12, hkv, bs, d, dtype=dtype, device=DEV)
    bt_dev = torch.tensor(bt, dtype=torch.int32, device=DEV)
    start = 0
    for n in (40, 1, 70, 39):
        slots = P.compute_slot_mapping_device(bt_dev, start, n, bs)
        P.reshape_and_cache(k[start:start + n].to(DEV), v[start:start + n].to(DEV), kv_dev, slots)
        got = P.paged_attention_prefill_chunk(q[start:start + n].to(DEV), kv_dev, bt_dev, start)
        ref = O.chunked_prefill_hbsd(q[start:start + n], kv_ref, bt, start, bs, compute_dtype=torch.float64)
        assert (got.double().cpu() - ref).abs().max().item() <= TOL[dtype]
        start += n
    assert torch.equal(kv_dev.cpu(), kv_ref)


def test_cpu_tensors_raise(P):
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        P.KVCachePool(4, 16, 2, 32, 1)                                     # reference default device='cpu'
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        P.paged_attention_decode(torch.randn(4, 32), torch.randn(2, 4, 4, 16, 32), [0], 3, 32, 4, 4, 16)
    kv = torch.zeros(2, 4, 4, 16, 32, device=DEV)
    with pytest.raises(IndexError):
        P.paged_attention_decode(torch.randn(4, 32, device=DEV), kv, [0], 40, 32, 4, 4, 16)   # table too short
    with pytest.raises(TypeError):
        P.paged_attention_decode(torch.randn(4, 32, device=DEV).half(), kv.half(), [0], 3, 32, 4, 4, 16)


def test_abi_status_codes(lib):
    from vllm_from_scratch_b200._lib import PaKvCache
    import ctypes
    from vllm_from_scratch_b200 import _lib as L
    assert lib.pa_abi_version() == L.ABI_VERSION == 3
    assert lib.pa_decode_num_workers() > 0
    c = PaKvCache(1, 1, 1, 1, 1, 1, 3, 16, 32, 0, 0)
    # Hq not a multiple of Hkv -> bad argument; never throws across the ABI
    rc = lib.pa_decode(16, 1, 1, 16, 1, 1, ctypes.byref(c), 16, 1, 16, 1, 4, 16, 1 << 20, None)
    assert rc == -1 and lib.pa_status_string(rc) == b"bad argument"
    c = PaKvCache(16, 16, 1, 1, 1, 1, 2, 16, 32, 7, 0)
    assert lib.pa_decode(16, 1, 1, 16, 1, 1, ctypes.byref(c), 16, 1, 16, 1, 4, 16, 1 << 20, None) == -2
    c = PaKvCache(16, 16, 512, 256, 32, 1, 2, 16, 32, 1, 0)
    assert lib.pa_decode(16, 1, 1, 16, 1, 1, ctypes.byref(c), 16, 1, 16, 1, 4, 16, 64, None) == -4   # workspace


# =========================================================================================
# fused RoPE + QKV split + KV slot write (mini_vllm.py:410-427) - bit-exact vs torch's op-by-op result
# =========================================================================================
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("hq,hkv,d,bs", [(8, 8, 32, 16), (4, 4, 16, 16), (32, 8, 128, 16), (2, 2, 6, 4), (8, 2, 64, 8)])
def test_rope_and_cache_bit_exact(P, dtype, hq, hkv, d, bs):
    g = torch.Generator().manual_seed(hq * 100 + d)
    t, max_len = 97, 512
    nb = (t + bs - 1) // bs + 3
    qkv = torch.randn(t, (hq + 2 * hkv) * d, generator=g).to(dtype)
    positions = torch.randint(0, max_len, (t,), generator=g)
    slots = torch.randperm(nb * bs, generator=g)[:t]
    slots[5] = -1                                                    # padding token: Q still rotates, no cache write
    cos, sin = (x.to(dtype) for x in O.rope_tables(d, max_len))     # nn.Module.to(dtype) casts the buffers too
    k_ref = torch.randn(nb, bs, hkv, d, generator=g).to(dtype)
    v_ref = torch.randn(nb, bs, hkv, d, generator=g).to(dtype)
    k_dev, v_dev = k_ref.to(DEV), v_ref.to(DEV)
    q_ref = O.rope_split_and_cache_bshd(qkv, positions, cos, sin, k_ref, v_ref, slots, hq, hkv)
    q_dev = P.rope_and_cache(qkv.to(DEV), positions.to(DEV), cos.to(DEV), sin.to(DEV), (k_dev, v_dev),
                             slots.to(DEV), hq)
    assert torch.equal(q_dev.cpu(), q_ref)
    assert torch.equal(k_dev.cpu(), k_ref) and torch.equal(v_dev.cpu(), v_ref)


def test_rope_and_cache_hbsd_layout_and_strided_qkv(P):
    """Module-01 pool layout [2, NB, Hkv, BS, D] and a qkv that is a column slice of a wider buffer."""
    g = torch.Generator().manual_seed(3)
    t, hq, hkv, d, bs, nb, max_len = 50, 8, 2, 64, 16, 6, 128
    wide = torch.randn(t, (hq + 2 * hkv) * d + 64, generator=g).to(torch.bfloat16)
    qkv = wide[:, 64:]
    positions = torch.arange(t)
    slots = torch.randperm(nb * bs, generator=g)[:t]
    cos, sin = (x.to(torch.bfloat16) for x in O.rope_tables(d, max_len))
    k_ref = torch.zeros(nb, bs, hkv, d, dtype=torch.bfloat16)
    v_ref = torch.zeros(nb, bs, hkv, d, dtype=torch.bfloat16)
    q_ref = O.rope_split_and_cache_bshd(qkv, positions, cos, sin, k_ref, v_ref, slots, hq, hkv)
    kv = torch.zeros(2, nb, hkv, bs, d, dtype=torch.bfloat16, device=DEV)
    q_dev = P.rope_and_cache(wide.to(DEV)[:, 64:], positions.to(DEV), cos.to(DEV), sin.to(DEV), kv, slots.to(DEV), hq)
    assert torch.equal(q_dev.cpu(), q_ref)
    assert torch.equal(kv[0].cpu().permute(0, 2, 1, 3), k_ref) and torch.equal(kv[1].cpu().permute(0, 2, 1, 3), v_ref)


@pytest.mark.parametrize("case", ROPE_CASES, ids=lambda c: c[0])
def test_rope_and_cache_reference_golden(P, case):
    """CUDA kernel vs the outputs of the reference's own RotaryEmbedding / rotate_half / scatter."""
    name, dt, h, d, bs, nb, t, max_len, seed = case
    gold = np.load(os.path.join(GOLD, "rope_cases.npz"))
    qkv, positions, slots, k_cache, v_cache = make_rope_inputs(case)
    assert tensor_sha(qkv, positions, slots, k_cache, v_cache) == str(gold[f"{name}/in_sha"]), "RNG drift"
    cos, sin = (x.to(qkv.dtype) for x in O.rope_tables(d, max_len))
    k_dev, v_dev = k_cache.to(DEV), v_cache.to(DEV)
    q = P.rope_and_cache(qkv.to(DEV), positions.to(DEV), cos.to(DEV), sin.to(DEV), (k_dev, v_dev), slots.to(DEV), h)
    assert torch.equal(q.float().cpu(), torch.from_numpy(gold[f"{name}/q"]))
    assert torch.equal(k_dev.float().cpu(), torch.from_numpy(gold[f"{name}/k_cache"]))
    assert torch.equal(v_dev.float().cpu(), torch.from_numpy(gold[f"{name}/v_cache"]))
