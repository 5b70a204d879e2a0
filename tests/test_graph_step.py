"""graph.DecodeStep: the captured step (H2D -> slot write -> decode -> D2H) against the oracle,
with inputs changed between replays (a decoding loop that appends one token per step)."""
import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import paged_attention_oracle as O  # noqa: E402

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype,tol", [(torch.bfloat16, 2e-2), (torch.float32, 1e-5)])
@pytest.mark.parametrize("capture", [True, False], ids=["graph", "eager"])
@pytest.mark.parametrize("zero_copy", [True, False], ids=["zerocopy", "staged"])
def test_decode_step_loop_matches_oracle(dtype, tol, capture, zero_copy):
    from vllm_from_scratch_b200 import paged_attention as P
    from vllm_from_scratch_b200.graph import DecodeStep
    dev = "cuda:0"
    g = torch.Generator().manual_seed(3)
    b, hq, hkv, d, bs, nb, maxb = 6, 8, 2, 128, 16, 64, 8
    kv_ref = torch.randn(2, nb, hkv, bs, d, generator=g).to(dtype)
    kv_dev = kv_ref.to(dev)
    perm = torch.randperm(nb, generator=g).to(torch.int32)
    bts = perm[:b * maxb].view(b, maxb)
    lens = torch.tensor([5, 16, 31, 47, 64, 100], dtype=torch.int32)
    step = DecodeStep(kv_dev, b, hq, maxb, zero_copy=zero_copy)
    if capture:
        step.capture()
    step.host["block_tables"].copy_(bts)
    for it in range(4):
        q = torch.randn(b, hq, d, generator=g).to(dtype)
        k = torch.randn(b, hkv, d, generator=g).to(dtype)
        v = torch.randn(b, hkv, d, generator=g).to(dtype)
        pos = lens.long()                                   # the new token's position
        slots = bts.long()[torch.arange(b), pos // bs] * bs + pos % bs
        lens = lens + 1
        step.host["q"].copy_(q)
        step.host["k"].copy_(k)
        step.host["v"].copy_(v)
        step.host["slots"].copy_(slots)
        step.host["seq_lens"].copy_(lens)
        step.replay()
        out = step.wait().clone()
        O.reshape_and_cache_hbsd(kv_ref, k, v, slots)
        ref = O.decode_batched(q, kv_ref, bts, lens, bs, compute_dtype=torch.float64)
        assert (out.double() - ref).abs().max().item() <= tol
        assert torch.equal(kv_dev.cpu(), kv_ref)            # the slot write is bit-exact
