"""Decode parity with bounds that actually constrain the long-context cases.

The north-star tolerance (max-abs 2e-2 in bf16) is loose for long contexts with N(0, 1) inputs: the
output of a 2k-token softmax average has std ~0.04 and at 32k ~0.01, so dropping a whole KV block - or
a whole split-K partial - stays below 2e-2.  Here the CUDA path is compared with the fp64 oracle under

    |out - ref| <= 2e-3 + 2^-7 * |ref|            (bf16: two roundings of 2^-9 relative, P and the output)
    |out - ref| <= 1e-5 * max(1, |ref|)           (fp32 KV cache: the north-star 1e-5)

on two kinds of inputs: N(0, 1), and PEAKED ones (every query head is a copy of one K row of its
sequence, placed in the first block, the last block and in between, so the softmax concentrates on
a few known tokens and the outputs are O(1) - a block or a partial that goes missing moves the result
by O(1)).  Every comparison comes with a NEGATIVE CONTROL: the same check must FAIL when seq_lens is
shortened by one block (and by one token where the peak is the last token).

fp32 at length (north_star "1e-5 with an fp32 KV cache", SURVEY.md section 7 "verify <= 1e-5 at ctx
2k-32k"): contexts 2048 / 8192 / 32768 against the fp64 oracle, and one sequence at 2048 against the
reference's own row-by-row fp32 arithmetic (oracle.decode_loop = paged_attention.py:80-101).
"""
import pytest
import torch

from oracle import paged_attention_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda"
BS = 16


@pytest.fixture(scope="module")
def P():
    from vllm_from_scratch_b200 import _lib, paged_attention as mod
    _lib.load()
    return mod


def _case(b, lens, hq, hkv, d, dtype, seed, peaked):
    """KV pool, scattered block tables and queries on the GPU (+ the host copies the oracle needs)."""
    g = torch.Generator(device=DEV).manual_seed(seed)
    gc = torch.Generator().manual_seed(seed)
    lens = torch.as_tensor(lens, dtype=torch.int32)
    nblk = (lens + BS - 1) // BS
    nb = int(nblk.sum()) + 8
    kv = torch.randn(2, nb, hkv, BS, d, generator=g, device=DEV, dtype=torch.float32).to(dtype)
    q = torch.randn(b, hq, d, generator=g, device=DEV, dtype=torch.float32).to(dtype)
    perm = torch.randperm(nb, generator=gc).to(torch.int32)
    bts = torch.zeros(b, int(nblk.max()), dtype=torch.int32)
    at = 0
    for i in range(b):
        n = int(nblk[i])
        bts[i, :n] = perm[at:at + n]
        at += n
    if peaked:
        grp = hq // hkv
        for i in range(b):
            n = int(lens[i])
            for h in range(hq):
                # head 0 of every group peaks on the LAST token, head 1 on the first block, the rest anywhere
                j = h % grp
                t = n - 1 if j == 0 else (int(torch.randint(0, min(BS, n), (1,), generator=gc)) if j == 1 else
                                         int(torch.randint(0, n, (1,), generator=gc)))
                q[i, h] = kv[0, int(bts[i, t // BS]), h // grp, t % BS]
    return q, kv, bts, lens


def _bound(ref, dtype):
    if dtype == torch.float32:
        return 1e-5 * ref.abs().clamp(min=1.0)
    return 2e-3 + 2.0 ** -7 * ref.abs()


def _violations(out, ref, dtype):
    """Worst ratio error / bound over all elements (<= 1 passes)."""
    return float(((out.double().cpu() - ref).abs() / _bound(ref, dtype)).max())


def _oracle(q, kv, bts, lens, rows=None):
    rows = list(range(q.shape[0])) if rows is None else rows
    q_c, kv_c = q.cpu(), kv.cpu()
    hq, d, hkv = q.shape[1], q.shape[2], kv.shape[2]
    return torch.stack([O.decode_vectorised(q_c[i], kv_c, bts[i].tolist(), int(lens[i]), d, hq, hkv, BS,
                                            compute_dtype=torch.float64) for i in rows]), rows


def _check_with_negative_controls(P, q, kv, bts, lens, dtype, rows=None, peaked=False):
    out = P.paged_attention_decode_batched(q, kv, bts.to(DEV), lens.to(DEV))
    ref, rows = _oracle(q, kv, bts, lens, rows)
    worst = _violations(out[rows], ref, dtype)
    assert worst <= 1.0, f"error / bound = {worst:.3f}"
    # negative control 1: one KV block less per sequence must NOT pass (N(0, 1) inputs in bf16 beyond 8k
    # tokens are excused: one block of 2048 moves a 32k-token average by less than a bf16 rounding - that is
    # what the peaked inputs are for)
    if peaked or dtype == torch.float32 or int(lens.max()) <= 8192:
        short = (lens - BS).clamp(min=1)
        out_short = P.paged_attention_decode_batched(q, kv, bts.to(DEV), short.to(DEV))
        assert _violations(out_short[rows], ref, dtype) > 1.0, "dropping a KV block went unnoticed: the bound is too loose"
    if peaked:
        # negative control 2: one TOKEN less - the heads that peak on the last token must move by O(1)
        out_m1 = P.paged_attention_decode_batched(q, kv, bts.to(DEV), (lens - 1).clamp(min=1).to(DEV))
        assert _violations(out_m1[rows], ref, dtype) > 10.0
    return out, worst


# ---------------------------------------------------------------------------------------
# bf16, the TMA split-K kernel, BASELINE.json shapes
# ---------------------------------------------------------------------------------------
@pytest.mark.parametrize("peaked", [False, True], ids=["normal", "peaked"])
@pytest.mark.parametrize("ragged", [False, True], ids=["equal", "ragged"])
def test_c2_64x2k_scale_aware(P, peaked, ragged):
    gc = torch.Generator().manual_seed(7)
    lens = torch.randint(1024, 2049, (64,), generator=gc).tolist() if ragged else [2048] * 64
    q, kv, bts, lens = _case(64, lens, 32, 8, 128, torch.bfloat16, seed=100 + peaked, peaked=peaked)
    out, _ = _check_with_negative_controls(P, q, kv, bts, lens, torch.bfloat16, peaked=peaked)
    if peaked:
        assert out.float().abs().max().item() > 1.0            # the outputs really are O(1)


@pytest.mark.parametrize("peaked", [False, True], ids=["normal", "peaked"])
def test_c3_8x32k_scale_aware(P, peaked):
    q, kv, bts, lens = _case(8, [32768] * 8, 32, 8, 128, torch.bfloat16, seed=110 + peaked, peaked=peaked)
    _check_with_negative_controls(P, q, kv, bts, lens, torch.bfloat16, rows=[0, 3, 7], peaked=peaked)


def test_c5_shard_64x8k_scale_aware(P):
    gc = torch.Generator().manual_seed(8)
    lens = torch.randint(8192 - 300, 8193, (64,), generator=gc).tolist()
    q, kv, bts, lens = _case(64, lens, 32, 8, 128, torch.bfloat16, seed=120, peaked=True)
    _check_with_negative_controls(P, q, kv, bts, lens, torch.bfloat16, rows=[0, 17, 40, 63], peaked=True)


@pytest.mark.parametrize("hkv", [1, 2])
def test_head_shard_rank_shapes_scale_aware(P, hkv):
    # per-rank shapes of KV-head-sharded 32k decode: few very long segments, up to 32 partials each
    q, kv, bts, lens = _case(4, [32768, 32768 - 77, 20000, 32768], 4 * hkv, hkv, 128, torch.bfloat16,
                             seed=130 + hkv, peaked=True)
    _check_with_negative_controls(P, q, kv, bts, lens, torch.bfloat16, peaked=True)


@pytest.mark.timeout(180)
def test_two_level_fold_uneven_partial_counts_complete_and_match(P):
    """8 x 32k with 4 KV heads (the per-rank shape of head-sharded C3 at 2 GPUs): the guided split gives segments of
    15, 16, 20, 24, 29 and 32 partials, i.e. two-level folds with ragged last groups, and one CTA can own several
    groups of one segment.  An early round-2 kernel deadlocked exactly here (the final fold queued in front of a
    group fold it polled for); the timeout turns a regression into a failure instead of a hung box.  Checked on every
    sequence with peaked inputs, negative controls included, and bit-reproducible across two launches."""
    q, kv, bts, lens = _case(8, [32768] * 8, 16, 4, 128, torch.bfloat16, seed=150, peaked=True)
    out, _ = _check_with_negative_controls(P, q, kv, bts, lens, torch.bfloat16, rows=[0, 1, 2, 4, 5, 7], peaked=True)
    again = P.paged_attention_decode_batched(q, kv, bts.to(DEV), lens.to(DEV))
    assert torch.equal(out, again)


def test_full_c2_batch_against_reference_row_by_row_arithmetic(P):
    """The whole 2048-token context of two C2 sequences against the reference's own bf16 row-by-row
    algorithm (oracle.decode_loop, bit-identical to paged_attention.py:80-101) - not a truncated prefix."""
    q, kv, bts, lens = _case(64, [2048] * 64, 32, 8, 128, torch.bfloat16, seed=140, peaked=False)
    out = P.paged_attention_decode_batched(q, kv, bts.to(DEV), lens.to(DEV))
    q_c, kv_c = q.cpu(), kv.cpu()
    for i in (5, 63):
        loop = O.decode_loop(q_c[i], kv_c, bts[i].tolist(), 2048, 128, 32, 8, BS)
        assert (out[i].float().cpu() - loop.float()).abs().max().item() <= 2e-2


# ---------------------------------------------------------------------------------------
# fp32 KV cache at length (generic kernel + reduce)
# ---------------------------------------------------------------------------------------
@pytest.mark.parametrize("peaked", [False, True], ids=["normal", "peaked"])
@pytest.mark.parametrize("ctx,b", [(2048, 6), (8192, 3), (32768, 2)])
def test_fp32_long_context_1e5(P, ctx, b, peaked):
    lens = [ctx] + [ctx - 1 - 37 * i for i in range(1, b)]
    q, kv, bts, lens = _case(b, lens, 32, 8, 128, torch.float32, seed=200 + ctx % 97 + peaked, peaked=peaked)
    _check_with_negative_controls(P, q, kv, bts, lens, torch.float32, peaked=peaked)


def test_fp32_2k_against_reference_fp32_loop(P):
    q, kv, bts, lens = _case(2, [2048, 1999], 32, 8, 128, torch.float32, seed=210, peaked=False)
    out = P.paged_attention_decode_batched(q, kv, bts.to(DEV), lens.to(DEV))
    loop = O.decode_loop(q[1].cpu(), kv.cpu(), bts[1].tolist(), 1999, 128, 32, 8, BS)     # reference arithmetic, fp32
    assert loop.dtype == torch.float32
    assert (out[1].cpu() - loop).abs().max().item() <= 1e-5


def test_fp32_engine_layout_d32_long(P):
    """The engine's own shape (head_dim 32, no GQA, 'bshd' buffers, mini_vllm.py:544-552) at 4k context."""
    g = torch.Generator(device=DEV).manual_seed(5)
    h, d, nblk = 8, 32, 256
    kc = torch.randn(2 * nblk + 4, BS, h, d, generator=g, device=DEV)
    vc = torch.randn(2 * nblk + 4, BS, h, d, generator=g, device=DEV)
    q = torch.randn(2, h, d, generator=g, device=DEV)
    perm = torch.randperm(2 * nblk + 4, generator=torch.Generator().manual_seed(5)).to(torch.int32)
    bts = perm[:2 * nblk].view(2, nblk)
    lens = torch.tensor([4096, 4001], dtype=torch.int32)
    out = P.paged_attention_decode_batched(q, (kc, vc), bts.to(DEV), lens.to(DEV))
    ref = O.decode_batched_bshd(q.cpu(), kc.cpu(), vc.cpu(), bts, lens, compute_dtype=torch.float64)
    assert _violations(out, ref, torch.float32) <= 1.0
