"""Seeded input builders shared by make_golden.py (run against the real reference in the
build container) and by the tests (CPU oracle tests and GPU parity tests).

Golden files store the reference OUTPUTS plus a sha256 of every regenerated INPUT, so a
drift in torch's CPU RNG would fail loudly instead of silently comparing different data.
"""
import hashlib

import numpy as np
import torch

# (name, dtype, Hq, Hkv, D, BS, NB, ctx, seed)
DECODE_CASES = [
    ("ref_test_equiv", "f32", 4, 4, 32, 4, 4, 8, 42),       # test_paged_attention.py:67-100 shape
    ("gqa_8_4_partial", "f32", 8, 4, 32, 16, 8, 20, 1),     # :59-65 shape
    ("gqa_8_2_d16", "f32", 8, 2, 16, 16, 2, 4, 2),          # :102-110 shape
    ("blog_bs2", "f32", 2, 2, 16, 2, 8, 6, 0),              # BLOG_PART1.md:629-663 shape
    ("ctx1", "f32", 4, 2, 32, 16, 4, 1, 3),
    ("ctx33_bs16", "f32", 8, 2, 64, 16, 8, 33, 4),
    ("g1_d128", "f32", 4, 4, 128, 16, 8, 50, 5),
    ("g4_d128", "f32", 32, 8, 128, 16, 24, 300, 6),         # Llama-3-8B head shape
    ("g8_d64_bs8", "f32", 16, 2, 64, 8, 16, 77, 7),
    ("g4_d128_bf16", "bf16", 32, 8, 128, 16, 24, 257, 8),
    ("g1_d64_bf16", "bf16", 4, 4, 64, 16, 8, 100, 9),
    ("g8_d128_bf16", "bf16", 16, 2, 128, 16, 12, 130, 10),
    ("g2_d32_bf16_bs4", "bf16", 8, 4, 32, 4, 16, 30, 11),
    ("g4_d128_bf16_exact_blocks", "bf16", 32, 8, 128, 16, 10, 128, 12),
    ("g4_d128_bf16_long", "bf16", 32, 8, 128, 16, 70, 1100, 13),
    ("g16_d128_bf16", "bf16", 16, 1, 128, 16, 12, 150, 14),
]

# (name, dtype, Hq, Hkv, D, BS, NB, T, seed)
PREFILL_CASES = [
    ("ref_test_written", "f32", 4, 4, 16, 4, 2, 4, 20),     # test_paged_attention.py:114-127 shape
    ("ref_test_shape", "f32", 4, 4, 32, 4, 4, 8, 21),       # :129-137 shape
    ("gqa_partial", "f32", 8, 2, 32, 16, 6, 37, 22),
    ("llama_bf16", "bf16", 32, 8, 128, 16, 6, 70, 23),
]


def _dtype(tag):
    return torch.float32 if tag == "f32" else torch.bfloat16


def tensor_sha(*tensors) -> str:
    h = hashlib.sha256()
    for t in tensors:
        if isinstance(t, torch.Tensor):
            t = t.detach().cpu().contiguous()
            if t.dtype == torch.bfloat16:
                t = t.view(torch.int16)
            h.update(t.numpy().tobytes())
        else:
            h.update(np.asarray(t).tobytes())
    return h.hexdigest()


def make_decode_inputs(case):
    name, dt, hq, hkv, d, bs, nb, ctx, seed = case
    g = torch.Generator().manual_seed(seed)
    kv = torch.randn(2, nb, hkv, bs, d, generator=g).to(_dtype(dt))
    q = torch.randn(hq, d, generator=g).to(_dtype(dt))
    nblk = (ctx + bs - 1) // bs
    bt = torch.randperm(nb, generator=g)[:nblk].tolist()
    return q, kv, bt


def make_prefill_inputs(case):
    name, dt, hq, hkv, d, bs, nb, t, seed = case
    g = torch.Generator().manual_seed(seed)
    q = torch.randn(t, hq, d, generator=g).to(_dtype(dt))
    k = torch.randn(t, hkv, d, generator=g).to(_dtype(dt))
    v = torch.randn(t, hkv, d, generator=g).to(_dtype(dt))
    nblk = (t + bs - 1) // bs
    bt = torch.randperm(nb, generator=g)[:nblk].tolist()
    kv = torch.zeros(2, nb, hkv, bs, d, dtype=_dtype(dt))
    return q, k, v, kv, bt


# (name, dtype, H, D, BS, NB, T, max_len, seed) - the engine's SelfAttention has Hq == Hkv == H
ROPE_CASES = [
    ("engine_default", "f32", 8, 32, 16, 8, 40, 256, 30),     # mini_vllm.py:943-954 head shape
    ("tiny_d16", "f32", 4, 16, 16, 4, 19, 64, 31),
    ("bf16_d128", "bf16", 2, 128, 16, 6, 70, 512, 32),
    ("bf16_d32", "bf16", 8, 32, 4, 32, 33, 128, 33),
]


def make_rope_inputs(case):
    name, dt, h, d, bs, nb, t, max_len, seed = case
    g = torch.Generator().manual_seed(seed)
    qkv = torch.randn(t, 3 * h * d, generator=g).to(_dtype(dt))
    positions = torch.randint(0, max_len, (t,), generator=g)
    slots = torch.randperm(nb * bs, generator=g)[:t]
    k_cache = torch.randn(nb, bs, h, d, generator=g).to(_dtype(dt))
    v_cache = torch.randn(nb, bs, h, d, generator=g).to(_dtype(dt))
    return qkv, positions, slots, k_cache, v_cache
