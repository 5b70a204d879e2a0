"""Row a8 / config C4: the REAL reference engine's step trace replayed on the CUDA kernels.

Expected side: ``tests/golden/engine_<case>.npz``, recorded by ``tests/golden/make_engine_golden.py`` from the
UNMODIFIED ``/root/reference/05_mini_vllm/mini_vllm.py`` (``MiniVLLM.step`` :712-764 driving the real
``Scheduler`` / ``BlockAllocator`` / ``MiniTransformer`` with ``engine_adapter`` + ``kv_manager`` over the CPU
oracle): 43 steps of chunked prefill (token budget 64) mixed with decode, 8 prompts sharing a 2-block prefix
through the prefix cache, 8 preemptions by swap in a 28-block pool.

Replay side (``tests/engine_replica.py``): the same scheduled requests go through the same
``engine_adapter.install`` wrapper into a replica of the model that holds bit-identical weights; the recorded
swap traffic goes through ``swap_out`` / ``swap_in``.  CPU test: oracle ops, proves the replica reproduces the
real engine's logits (so any GPU mismatch is the kernels').  GPU tests: this package's CUDA ops, logits compared
step by step - fp32 within 1e-4 and identical greedy tokens; the bf16 / head_dim 128 case (TMA decode, tcgen05
varlen prefill) within the bf16 tolerance.
"""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "tests", "golden")):
    if p not in sys.path:
        sys.path.insert(0, p)

from engine_cases import ENGINE_CASES, det_weights, torch_dtype  # noqa: E402
from engine_harness import OracleOps  # noqa: E402
from engine_replica import ReplayEngine, Trace  # noqa: E402
from make_engine_golden import weights_sha  # noqa: E402
from vllm_from_scratch_b200 import engine_adapter  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")


class _CpuOps(OracleOps):
    """Oracle ops with the swap return convention of the CUDA module (an event or None)."""


def _replay(case, device, ops):
    cfg = ENGINE_CASES[case]
    trace = Trace(os.path.join(GOLDEN, f"engine_{case}.npz"))
    weights = det_weights(cfg)
    assert weights_sha(weights) == trace.weights_sha, "seeded weights differ from the ones the golden run used"
    if ops is None:
        from vllm_from_scratch_b200 import paged_attention as ops      # the product path: CUDA kernels, no fallback
    eng = ReplayEngine(cfg, weights, trace, device, torch_dtype(cfg), ops)
    engine_adapter.install(eng, ops=ops)
    stats = {"max_err": 0.0, "rows": 0, "tok_same": 0, "tok_checked": 0, "full_rows": 0, "scale": 0.0, "steps": 0}
    for s in range(trace.num_steps):
        logits = eng.step().float().cpu().numpy()
        nxt, tv, ti, full = trace.expected(s)
        assert logits.shape[0] == len(nxt)
        assert np.isfinite(logits).all(), f"step {s}: non-finite logits (a poisoned block was read)"
        got_top = np.take_along_axis(logits, ti, axis=1)
        stats["max_err"] = max(stats["max_err"], float(np.abs(got_top - tv).max()))
        if full is not None:
            stats["max_err"] = max(stats["max_err"], float(np.abs(logits - full).max()))
            stats["full_rows"] += full.shape[0]
        stats["scale"] = max(stats["scale"], float(np.abs(tv).max()))
        margin = tv[:, 0] - tv[:, 1]
        sure = margin > 1e-3 * max(1.0, float(np.abs(tv).max()))
        stats["tok_checked"] += int(sure.sum())
        stats["tok_same"] += int((logits.argmax(1)[sure] == nxt[sure]).sum())
        stats["rows"] += len(nxt)
        stats["steps"] += 1
    stats["swapped_blocks"] = eng.swapped_blocks
    return stats, trace


def test_trace_exercises_sharing_swap_and_mixed_steps():
    for case in ENGINE_CASES:
        z = np.load(os.path.join(GOLDEN, f"engine_{case}.npz"))
        hit_blocks, _, preemptions, out_blocks, in_blocks = z["stats"].tolist()
        assert hit_blocks > 0 and preemptions > 0 and out_blocks == in_blocks > 0
        mixed = 0
        for s in range(len(z["step_ptr"]) - 1):
            pf = z["is_prefill"][z["step_ptr"][s]:z["step_ptr"][s + 1]]
            mixed += int(pf.any() and not pf.all())
        assert mixed >= 5
        # integer contract of the recorded scheduler: slot = block_table[pos // BS] * BS + pos % BS (mini_vllm.py:209-219)
        bs = ENGINE_CASES[case]["block_size"]
        for r in range(len(z["req_ids"])):
            bt = z["bt_flat"][z["bt_ptr"][r]:z["bt_ptr"][r + 1]]
            n = z["tok_ptr"][r + 1] - z["tok_ptr"][r]
            pos = z["start_pos"][r] + np.arange(n)
            assert (z["slots_flat"][z["tok_ptr"][r]:z["tok_ptr"][r + 1]] == bt[pos // bs] * bs + pos % bs).all()


@pytest.mark.parametrize("case", list(ENGINE_CASES))
def test_replica_with_oracle_ops_reproduces_real_engine_logits_cpu(case):
    stats, trace = _replay(case, "cpu", _CpuOps)
    tol = 2e-5 if ENGINE_CASES[case]["dtype"] == "f32" else 1e-6      # bf16 on one CPU: the same arithmetic, bit for bit
    assert stats["steps"] == trace.num_steps and stats["full_rows"] > 0 and stats["swapped_blocks"] > 0
    assert stats["max_err"] <= tol * max(1.0, stats["scale"]), stats
    assert stats["tok_same"] == stats["tok_checked"] > 0.9 * stats["rows"], stats


@pytest.mark.gpu
def test_real_engine_trace_on_cuda_kernels_f32():
    """C4 as the reference ships it (fp32, head_dim 32): generic decode / prefill kernels, fused RoPE + slot
    write, block swap - logits of every step vs the real engine's, <= 1e-4 (relative to the logit scale), and
    the same greedy tokens."""
    stats, trace = _replay("c4_f32", "cuda", None)
    assert stats["steps"] == trace.num_steps and stats["full_rows"] > 0 and stats["swapped_blocks"] > 0
    assert stats["max_err"] <= 1e-4 * max(1.0, stats["scale"]), stats
    assert stats["tok_same"] == stats["tok_checked"] > 0.9 * stats["rows"], stats


@pytest.mark.gpu
def test_real_engine_trace_on_cuda_kernels_bf16_d128():
    """The same loop in bf16 with head_dim 128: every mixed step is one tcgen05 varlen launch, decode-only
    steps too; KV written by the fused RoPE kernel.  bf16 GEMMs differ between the CPU and cuBLAS, so the bound
    is the bf16 one: 5 % of the logit scale, and >= 90 % identical greedy tokens on rows with a clear margin."""
    from vllm_from_scratch_b200 import paged_attention as P
    cfg = ENGINE_CASES["c4_bf16_d128"]
    kc = torch.zeros(4, 16, cfg["num_heads"], 128, dtype=torch.bfloat16, device="cuda")
    assert P.prefill_tensor_core_available((kc, kc.clone()), cfg["num_heads"])
    stats, trace = _replay("c4_bf16_d128", "cuda", None)
    assert stats["steps"] == trace.num_steps and stats["swapped_blocks"] > 0
    assert stats["max_err"] <= 5e-2 * max(1.0, stats["scale"]), stats
    assert stats["tok_same"] >= 0.9 * stats["tok_checked"], stats
