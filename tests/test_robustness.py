"""Failure modes the happy-path parity tests do not reach: corrupted block ids, short block tables, more
than one GPU driven from one process (per-device kernel attributes), tensor-map cache turnover."""
import pytest
import torch

from oracle import paged_attention_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.fixture(scope="module")
def P():
    from vllm_from_scratch_b200 import _lib, paged_attention as mod
    _lib.load()
    return mod


def _batch(dev, nb=40, hq=32, hkv=8, d=128, bs=16, lens=(160, 145, 7), seed=2, dtype=torch.bfloat16):
    g = torch.Generator().manual_seed(seed)
    kv = torch.randn(2, nb, hkv, bs, d, generator=g).to(dtype).to(dev)
    q = torch.randn(len(lens), hq, d, generator=g).to(dtype).to(dev)
    perm = torch.randperm(nb, generator=g).to(torch.int32)
    nblk = [(n + bs - 1) // bs for n in lens]
    bts = torch.zeros(len(lens), max(nblk), dtype=torch.int32)
    at = 0
    for i, n in enumerate(nblk):
        bts[i, :n] = perm[at:at + n]
        at += n
    return q, kv, bts, torch.tensor(lens, dtype=torch.int32)


def test_block_id_outside_the_cache_reads_zeros_not_memory(P):
    """The TMA tensor map's block axis is bounded by the cache's own block count: a block id >= NB (a corrupted
    table) contributes zero K / V rows instead of reading whatever lies behind the cache."""
    q, kv, bts, lens = _batch(DEV)
    big = torch.empty(2, 80, 8, 16, 128, dtype=torch.bfloat16, device=DEV)
    big[:] = float("nan")                                       # memory "behind" the 40-block cache is poison
    big[:, :40] = kv
    cache = big[:, :40]                                         # the cache proper: NB = 40
    bad = bts.clone()
    bad[0, 3] = 41 + 5                                          # out of range, lands in the poisoned tail
    out = P.paged_attention_decode_batched(q, cache, bad.to(DEV), lens.to(DEV))
    assert torch.isfinite(out.float()).all(), "an out-of-range block id read memory outside the cache"
    # the in-range sequences are untouched
    ref = P.paged_attention_decode_batched(q, cache, bts.to(DEV), lens.to(DEV))
    assert torch.equal(out[1:], ref[1:])
    # and sequence 0 equals the decode with that block's K = V = 0
    zeroed = kv.clone()
    spare = int(set(range(40)).difference(bts.flatten().tolist()).pop())
    zeroed[:, spare] = 0
    fix = bts.clone()
    fix[0, 3] = spare
    want = O.decode_batched(q.cpu(), zeroed.cpu(), fix, lens, 16, compute_dtype=torch.float64)
    assert (out[0].double().cpu() - want[0]).abs().max().item() <= 2e-2


def test_prefill_chunk_short_block_table_raises_like_the_reference(P):
    q, kv, bts, lens = _batch(DEV)
    qq = torch.randn(20, 32, 128, device=DEV).to(torch.bfloat16)
    with pytest.raises(IndexError):                             # positions 30..49 need 4 blocks, the table has 3
        P.paged_attention_prefill_chunk(qq, kv, bts[0, :3].to(DEV), 30)
    with pytest.raises(IndexError):                             # the reference's block_table[i // BS] (paged_attention.py:119-123)
        P.paged_attention_prefill(qq, qq[:, :8].contiguous(), qq[:, :8].contiguous(), kv, [1], 16)


def test_varlen_request_with_a_short_table_row_is_skipped_not_read_past(P):
    q, kv, bts, lens = _batch(DEV, lens=(64, 64))
    t = 8
    qq = torch.randn(2 * t, 32, 128, device=DEV).to(torch.bfloat16)
    cu = torch.tensor([0, t, 2 * t], dtype=torch.int32, device=DEV)
    ctx_ok = torch.tensor([64, 64], dtype=torch.int32, device=DEV)
    ctx_bad = torch.tensor([64, 5000], dtype=torch.int32, device=DEV)      # row 1 claims 313 blocks, the table has 4
    good = P.paged_attention_prefill_varlen(qq, kv, bts.to(DEV), cu, ctx_ok, t)
    out = torch.full_like(good, 7.0)
    P.paged_attention_prefill_varlen(qq, kv, bts.to(DEV), cu, ctx_bad, t, out=out)
    assert torch.equal(out[:t], good[:t])
    assert bool((out[t:] == 7.0).all())                                     # skipped: rows left alone


def test_tensor_map_cache_survives_a_32_layer_engine_working_set(P):
    """64 decode maps + 64 prefill maps (K and V of 32 layers, two box shapes) cycled three times: results stay
    right when the cache turns over (it is hashed LRU with 1024 entries; this is the functional check)."""
    layers = 32
    g = torch.Generator().manual_seed(3)
    pool = torch.randn(2, layers, 12, 8, 16, 128, generator=g).to(torch.bfloat16).to(DEV)
    q = torch.randn(2, 32, 128, generator=g).to(torch.bfloat16).to(DEV)
    bts = torch.tensor([[3, 1, 7, 0], [2, 9, 11, 5]], dtype=torch.int32)
    lens = torch.tensor([64, 50], dtype=torch.int32)
    cu = torch.tensor([0, 1, 2], dtype=torch.int32, device=DEV)
    first = None
    for rep in range(3):
        outs = []
        for layer in range(layers):
            cache = pool[:, layer]
            d = P.paged_attention_decode_batched(q, cache, bts.to(DEV), lens.to(DEV))
            v = P.paged_attention_prefill_varlen(q, cache, bts.to(DEV), cu, lens.to(DEV), 1)
            outs.append((d, v))
        if first is None:
            first = outs
            for layer in (0, 17, 31):
                ref = O.decode_batched(q.cpu(), pool[:, layer].cpu(), bts, lens, 16, compute_dtype=torch.float64)
                assert (outs[layer][0].double().cpu() - ref).abs().max().item() <= 2e-2
                assert (outs[layer][1].double().cpu() - ref).abs().max().item() <= 2e-2
        else:
            for (d0, v0), (d1, v1) in zip(first, outs):
                assert torch.equal(d0, d1) and torch.equal(v0, v1)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs in one process")
def test_two_devices_in_one_process(P):
    """cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is per device: the first launch of the 200 KB-smem
    kernels on a SECOND GPU of the same process must work (decode fast path, both prefill kernels, generic)."""
    from vllm_from_scratch_b200 import _lib
    lib = _lib.load()
    outs = []
    for dev in ("cuda:0", "cuda:1", "cuda:0"):
        q, kv, bts, lens = _batch(dev, lens=(300, 145, 7))
        o = P.paged_attention_decode_batched(q, kv, bts.to(dev), lens.to(dev))
        qq = torch.randn(40, 32, 128, generator=torch.Generator().manual_seed(1)).to(torch.bfloat16).to(dev)
        pf = P.paged_attention_prefill_chunk(qq, kv, bts[0].to(dev), 200)
        old = lib.pa_prefill_force_path(2)
        try:
            pf1 = P.paged_attention_prefill_chunk(qq, kv, bts[0].to(dev), 200)
        finally:
            lib.pa_prefill_force_path(old)
        qf, kvf, btf, lf = _batch(dev, dtype=torch.float32, lens=(100, 33))
        of = P.paged_attention_decode_batched(qf, kvf, btf.to(dev), lf.to(dev))
        torch.cuda.synchronize(dev)
        outs.append((o.cpu(), pf.cpu(), pf1.cpu(), of.cpu()))
        ref = O.decode_batched(q.cpu(), kv.cpu(), bts, lens, 16, compute_dtype=torch.float64)
        assert (o.double().cpu() - ref).abs().max().item() <= 2e-2
    for a, b in zip(outs[0], outs[1]):
        assert torch.equal(a, b)                                # same inputs, same bits on both GPUs
