"""GPU parity tests of the tensor-core (tcgen05) paged prefill kernel, csrc/pa_prefill_tc.cu,
called through the Python host -> C ABI (pa_prefill / pa_prefill_varlen).

Oracle: oracle.chunked_prefill_hbsd in fp64 on the same bf16 inputs (for start_pos == 0 it is the
reference's paged_attention_prefill, paged_attention.py:104-136, which tests/golden pins).
Tolerance: max-abs <= 2e-2 (BASELINE.json north_star, bf16).
"""
import pytest
import torch

from oracle import paged_attention_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda"
TOL = 2e-2


@pytest.fixture(scope="module")
def P():
    from vllm_from_scratch_b200 import _lib, paged_attention as mod
    _lib.load()
    return mod


@pytest.fixture(scope="module")
def lib():
    from vllm_from_scratch_b200 import _lib
    return _lib.load()


@pytest.fixture(autouse=True, params=[0, 2], ids=["pingpong", "onetile"])
def tc_kernel(request, lib):
    """Every test runs through both tcgen05 kernels: the two-tile ping-pong one (default) and the
    one-tile-per-CTA one (pa_prefill_force_path(2))."""
    old = lib.pa_prefill_force_path(request.param)
    yield request.param
    lib.pa_prefill_force_path(old)


def _make_seq(g, total, hq, hkv, d, bs, nb, poison=False):
    """One sequence of `total` tokens scattered over a pool of nb blocks."""
    q = torch.randn(total, hq, d, generator=g).to(torch.bfloat16)
    k = torch.randn(total, hkv, d, generator=g).to(torch.bfloat16)
    v = torch.randn(total, hkv, d, generator=g).to(torch.bfloat16)
    nblk = (total + bs - 1) // bs
    bt = torch.randperm(nb, generator=g)[:nblk].tolist()
    kv = torch.zeros(2, nb, hkv, bs, d, dtype=torch.bfloat16)
    if poison:
        kv.fill_(float("nan"))          # slots nobody wrote must never reach the output
    O.write_kv_hbsd(kv, bt, k, v, bs)
    return q, kv, bt


@pytest.mark.parametrize("hq,hkv,d", [(32, 8, 128), (8, 8, 128), (8, 4, 64), (16, 2, 128), (16, 1, 128), (4, 4, 64)])
@pytest.mark.parametrize("total", [1, 16, 70, 128, 129, 300, 1000])
def test_prefill_tc_full_sequence(P, hq, hkv, d, total):
    g = torch.Generator().manual_seed(1000 * hq + total)
    bs, nb = 16, 80
    q, kv, bt = _make_seq(g, total, hq, hkv, d, bs, nb, poison=True)
    assert P.prefill_tensor_core_available(kv.to(DEV), hq)
    bt_dev = torch.tensor(bt, dtype=torch.int32, device=DEV)
    got = P.paged_attention_prefill_chunk(q.to(DEV), kv.to(DEV), bt_dev, 0)
    ref = O.chunked_prefill_hbsd(q, torch.nan_to_num(kv), bt, 0, bs, compute_dtype=torch.float64)
    err = (got.double().cpu() - ref).abs().max().item()
    assert err <= TOL, err


@pytest.mark.parametrize("start,n", [(0, 40), (40, 1), (41, 200), (241, 15), (256, 128), (384, 131), (515, 700)])
def test_prefill_tc_chunk_with_history(P, start, n):
    g = torch.Generator().manual_seed(7)
    hq, hkv, d, bs, nb = 32, 8, 128, 16, 100
    q, kv, bt = _make_seq(g, 1215, hq, hkv, d, bs, nb)
    kv_dev = kv.to(DEV)
    bt_dev = torch.tensor(bt, dtype=torch.int32, device=DEV)
    got = P.paged_attention_prefill_chunk(q[start:start + n].to(DEV), kv_dev, bt_dev, start)
    ref = O.chunked_prefill_hbsd(q[start:start + n], kv, bt, start, bs, compute_dtype=torch.float64)
    err = (got.double().cpu() - ref).abs().max().item()
    assert err <= TOL, err


def test_prefill_tc_matches_decode_formulation(P, lib):
    """Same inputs through both CUDA implementations behind pa_prefill."""
    g = torch.Generator().manual_seed(11)
    hq, hkv, d, bs, nb = 32, 8, 128, 16, 40
    q, kv, bt = _make_seq(g, 333, hq, hkv, d, bs, nb)
    kv_dev, q_dev = kv.to(DEV), q.to(DEV)
    bt_dev = torch.tensor(bt, dtype=torch.int32, device=DEV)
    tc = P.paged_attention_prefill_chunk(q_dev[100:], kv_dev, bt_dev, 100)
    cur = lib.pa_prefill_force_path(1)
    try:
        dec = P.paged_attention_prefill_chunk(q_dev[100:], kv_dev, bt_dev, 100)
    finally:
        lib.pa_prefill_force_path(cur)
    assert (tc.float() - dec.float()).abs().max().item() <= TOL
    again = P.paged_attention_prefill_chunk(q_dev[100:], kv_dev, bt_dev, 100)
    assert torch.equal(tc, again)                     # bit-reproducible


def _varlen_batch(g, specs, hq, hkv, d, bs, nb, layout="hbsd"):
    """specs = [(ctx_len, q_len)].  Returns packed query, caches, tables, the per-sequence oracle output."""
    kv = torch.zeros(2, nb, hkv, bs, d, dtype=torch.bfloat16)
    perm = torch.randperm(nb, generator=g).tolist()
    maxblk = max((c + bs - 1) // bs for c, _ in specs)
    bts = torch.zeros(len(specs), maxblk, dtype=torch.int32)
    qs, refs, cu, at = [], [], [0], 0
    for i, (ctx, ql) in enumerate(specs):
        nblk = (ctx + bs - 1) // bs
        bt = perm[at:at + nblk]
        at += nblk
        bts[i, :nblk] = torch.tensor(bt, dtype=torch.int32)
        k = torch.randn(ctx, hkv, d, generator=g).to(torch.bfloat16)
        v = torch.randn(ctx, hkv, d, generator=g).to(torch.bfloat16)
        O.write_kv_hbsd(kv, bt, k, v, bs)
        q = torch.randn(ql, hq, d, generator=g).to(torch.bfloat16)
        qs.append(q)
        cu.append(cu[-1] + ql)
        if ql:
            refs.append(O.chunked_prefill_hbsd(q, kv, bt, ctx - ql, bs, compute_dtype=torch.float64))
    return (torch.cat(qs), kv, bts, torch.tensor(cu, dtype=torch.int32),
            torch.tensor([c for c, _ in specs], dtype=torch.int32), torch.cat(refs))


@pytest.mark.parametrize("hq,hkv,d", [(32, 8, 128), (8, 2, 64)])
def test_prefill_tc_varlen_mixed_batch(P, hq, hkv, d):
    g = torch.Generator().manual_seed(5)
    bs, nb = 16, 160
    specs = [(300, 300), (77, 1), (512, 64), (1, 1), (130, 0), (640, 129), (33, 33), (257, 1)]
    q, kv, bts, cu, ctx, ref = _varlen_batch(g, specs, hq, hkv, d, bs, nb)
    got = P.paged_attention_prefill_varlen(q.to(DEV), kv.to(DEV), bts.to(DEV), cu.to(DEV), ctx.to(DEV),
                                           max(ql for _, ql in specs))
    err = (got.double().cpu() - ref).abs().max().item()
    assert err <= TOL, err


def test_prefill_tc_varlen_equals_decode_for_single_tokens(P):
    """q_len == 1 everywhere: the varlen prefill IS batched decode."""
    g = torch.Generator().manual_seed(6)
    hq, hkv, d, bs, nb = 32, 8, 128, 16, 120
    specs = [(n, 1) for n in (1, 15, 16, 17, 200, 511, 640)]
    q, kv, bts, cu, ctx, ref = _varlen_batch(g, specs, hq, hkv, d, bs, nb)
    kv_dev = kv.to(DEV)
    got = P.paged_attention_prefill_varlen(q.to(DEV), kv_dev, bts.to(DEV), cu.to(DEV), ctx.to(DEV), 1)
    dec = P.paged_attention_decode_batched(q.to(DEV), kv_dev, bts.to(DEV), ctx.to(DEV))
    assert (got.double().cpu() - ref).abs().max().item() <= TOL
    assert (got.float() - dec.float()).abs().max().item() <= TOL


def test_prefill_tc_bshd_layout(P):
    """The engine layout (k_cache, v_cache)[NB, BS, H, D] through the same kernel (strides only)."""
    g = torch.Generator().manual_seed(8)
    hq, hkv, d, bs, nb, total = 8, 8, 64, 16, 30, 200
    q, kv, bt = _make_seq(g, total, hq, hkv, d, bs, nb)
    k_cache = kv[0].permute(0, 2, 1, 3).contiguous().to(DEV)       # [NB, BS, H, D]
    v_cache = kv[1].permute(0, 2, 1, 3).contiguous().to(DEV)
    bt_dev = torch.tensor(bt, dtype=torch.int32, device=DEV)
    got = P.paged_attention_prefill_chunk(q[50:].to(DEV), (k_cache, v_cache), bt_dev, 50)
    ref = O.chunked_prefill_hbsd(q[50:], kv, bt, 50, bs, compute_dtype=torch.float64)
    assert (got.double().cpu() - ref).abs().max().item() <= TOL


def test_prefill_tc_long_context_chunk(P):
    """A 256-token chunk at the end of a 16k context: 128 KV tiles per work tile."""
    g = torch.Generator().manual_seed(9)
    hq, hkv, d, bs = 32, 8, 128, 16
    total, n = 16384, 256
    nb = total // bs + 8
    q, kv, bt = _make_seq(g, total, hq, hkv, d, bs, nb)
    bt_dev = torch.tensor(bt, dtype=torch.int32, device=DEV)
    got = P.paged_attention_prefill_chunk(q[total - n:].to(DEV), kv.to(DEV), bt_dev, total - n)
    ref = O.chunked_prefill_hbsd(q[total - n:], kv, bt, total - n, bs, compute_dtype=torch.float64)
    assert (got.double().cpu() - ref).abs().max().item() <= TOL


def test_prefill_tc_persistent_scheduler_sparse_batch(P):
    """More work tiles than SMs, most of them empty: one long prompt next to many one-token rows and masked rows
    (q_len 0).  The persistent kernel's scheduler warp has to skip the empty tiles of the launch's ticket list,
    every CTA walks several tiles (barrier phases carried across tiles), and tiles without a second query tile
    (a warpgroup with no rows) sit between tiles that have one."""
    g = torch.Generator().manual_seed(21)
    hq, hkv, d, bs = 32, 8, 128, 16
    specs = [(1400, 1400)] + [(40 + 7 * i, 1) for i in range(40)] + [(64, 0)] * 6 + [(333, 97), (200, 33), (96, 96)]
    nb = sum((c + bs - 1) // bs for c, _ in specs) + 8
    q, kv, bts, cu, ctx, ref = _varlen_batch(g, specs, hq, hkv, d, bs, nb)
    args = (q.to(DEV), kv.to(DEV), bts.to(DEV), cu.to(DEV), ctx.to(DEV), max(ql for _, ql in specs))
    got = P.paged_attention_prefill_varlen(*args)
    assert (got.double().cpu() - ref).abs().max().item() <= TOL
    # the launch leaves its ticket slot clean: the same call again (and again) gives the same bits
    for _ in range(3):
        again = P.paged_attention_prefill_varlen(*args)
        assert torch.equal(again, got)


def test_prefill_tc_concurrent_launches_on_two_streams(P):
    """Two launches in flight at once draw their work tiles from different ticket slots."""
    g = torch.Generator().manual_seed(22)
    hq, hkv, d, bs = 32, 8, 128, 16
    batches = []
    for specs in ([(900, 900), (300, 1), (650, 130)], [(1100, 700), (80, 80), (500, 1), (64, 0)]):
        nb = sum((c + bs - 1) // bs for c, _ in specs) + 4
        q, kv, bts, cu, ctx, ref = _varlen_batch(g, specs, hq, hkv, d, bs, nb)
        batches.append(((q.to(DEV), kv.to(DEV), bts.to(DEV), cu.to(DEV), ctx.to(DEV), max(ql for _, ql in specs)), ref))
    torch.cuda.synchronize()
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    outs = [[], []]
    for _ in range(4):
        for i, (args, _) in enumerate(batches):
            with torch.cuda.stream(streams[i]):
                outs[i].append(P.paged_attention_prefill_varlen(*args))
    torch.cuda.synchronize()
    for i, (_, ref) in enumerate(batches):
        for got in outs[i]:
            assert (got.double().cpu() - ref).abs().max().item() <= TOL
