"""Prefix-cache wiring and preemption-by-swap in the engine loop.

CPU part (build container): ``kv_manager`` over the REAL reference engine
(/root/reference/05_mini_vllm/mini_vllm.py, unmodified; skipped where the reference is absent) and
over the TinyEngine harness, oracle-backed ops.  The property checked everywhere: sharing blocks
and moving a request's KV through the host changes NOTHING a request computes - same tokens, same
logits as an engine with a private, big-enough pool.
GPU part: the same scenarios on cuda through the CUDA swap / attention kernels.
"""
import json
import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from engine_harness import OracleOps, TinyEngine  # noqa: E402
from vllm_from_scratch_b200 import engine_adapter, kv_manager  # noqa: E402

REF_DIR = "/root/reference/05_mini_vllm"
needs_reference = pytest.mark.skipif(not os.path.exists(os.path.join(REF_DIR, "mini_vllm.py")),
                                     reason="the reference tree is only present in the build container")
GOLDEN = os.path.join(ROOT, "tests", "golden", "block_hash_cases.json")


def _ref_module():
    if REF_DIR not in sys.path:
        sys.path.insert(0, REF_DIR)
    import mini_vllm
    return mini_vllm


# ---------------------------------------------------------------------------------------
# chained block hashes: golden vectors produced by the reference (tests/golden/make_hash_golden.py)
# ---------------------------------------------------------------------------------------
def test_block_hashes_match_reference_golden():
    cases = json.load(open(GOLDEN))["cases"]
    assert len(cases) >= 6
    for c in cases:
        got = kv_manager.compute_block_hashes(c["token_ids"], c["block_size"], c["extra_key"])
        assert got == c["hashes"], c["name"]


def test_block_hashes_chain_properties():
    a = kv_manager.compute_block_hashes(list(range(40)), 16)
    assert len(a) == 2                                           # only FULL blocks
    b = kv_manager.compute_block_hashes(list(range(32)) + [99] * 8, 16)
    assert a == b                                                # the partial tail does not matter
    c = kv_manager.compute_block_hashes([7] + list(range(1, 32)), 16)
    assert c[0] != a[0] and c[1] != a[1]                         # a changed prefix changes every later hash
    d = kv_manager.compute_block_hashes(list(range(16)) + [5] + list(range(17, 32)), 16)
    assert d[0] == a[0] and d[1] != a[1]
    assert kv_manager.compute_block_hashes(list(range(15)), 16) == []
    assert kv_manager.compute_block_hashes(list(range(32)), 16, extra_key=3) != a


# ---------------------------------------------------------------------------------------
# TinyEngine scenarios (shared by the CPU and the GPU tests)
# ---------------------------------------------------------------------------------------
PREFIX = list(range(3, 3 + 37))                                  # 2 full blocks + 5 tokens
PROMPTS = [PREFIX + [100 + i, 120 + i, 140 + i] * (2 + i) for i in range(4)] + [list(range(60, 60 + 21))]


def _tiny(device, ops, num_blocks, manager, dtype=torch.float32, token_budget=48, max_new=10, stagger=True, **kw):
    eng = TinyEngine(device=device, dtype=dtype, num_blocks=num_blocks, token_budget=token_budget,
                     hashed_blocks=True, **kw)
    engine_adapter.install(eng, ops=ops)
    mgr = kv_manager.install(eng, num_host_blocks=64, ops=ops, **manager) if manager is not None else None
    reqs = [eng.add_request(PROMPTS[0], max_new_tokens=max_new)]
    pending = list(PROMPTS[1:])
    steps = 0
    while (eng.scheduler.has_unfinished or pending) and steps < 400:
        if pending and (not stagger or steps >= 2):
            reqs.append(eng.add_request(pending.pop(0), max_new_tokens=max_new))   # one new request per step
        eng.step()
        steps += 1
    assert not eng.scheduler.has_unfinished and not pending
    return eng, reqs, mgr


def _check_same(a, ra, b, rb, atol):
    for x, y in zip(ra, rb):
        assert x.output_token_ids == y.output_token_ids, (x.request_id, x.output_token_ids, y.output_token_ids)
        la, lb = a.logits_log[x.request_id], b.logits_log[y.request_id]
        assert len(la) == len(lb)
        for u, v in zip(la, lb):
            assert torch.allclose(u, v, atol=atol), (u - v).abs().max()


def test_tiny_prefix_cache_shares_blocks_and_changes_nothing_cpu():
    base, rb, _ = _tiny("cpu", OracleOps, 64, None)
    eng, rs, mgr = _tiny("cpu", OracleOps, 64, dict(enable_swap=False))
    _check_same(base, rb, eng, rs, 1e-4)
    # requests 1..3 arrive after request 0 filled the two prefix blocks: 2 blocks / 32 tokens each
    assert mgr.stats.prefix_hit_blocks == 6 and mgr.stats.prefix_hit_tokens == 96
    assert mgr.stats.blocks_marked > 0
    alloc = eng.blocks
    assert alloc.num_free == 64 and all(c == 0 for c in alloc.ref_count)     # every reference returned
    # a later identical prompt revives the freed-but-hashed blocks (all full blocks but the last token's)
    r = eng.add_request(PROMPTS[0], max_new_tokens=3)
    eng.run()
    assert r.output_token_ids == rs[0].output_token_ids[:3]
    assert mgr.stats.prefix_hit_blocks == 6 + (len(PROMPTS[0]) - 1) // 16


def test_tiny_prefix_hit_uses_the_donor_blocks_cpu():
    eng = TinyEngine(device="cpu", num_blocks=32, hashed_blocks=True, token_budget=256)
    engine_adapter.install(eng, ops=OracleOps)
    mgr = kv_manager.install(eng, ops=OracleOps, enable_swap=False)
    a = eng.add_request(PROMPTS[0], max_new_tokens=8)
    eng.step()                                                   # whole prompt in one chunk
    b = eng.add_request(PROMPTS[1], max_new_tokens=8)
    eng.step()
    assert b.block_table[:2] == a.block_table[:2]                # by reference, not by copy
    assert eng.blocks.ref_count[a.block_table[0]] == 2 and eng.blocks.ref_count[a.block_table[1]] == 2
    assert b.block_table[2] != a.block_table[2]                  # the partial third block is private
    assert mgr.stats.prefix_hit_tokens == 32
    eng.run()
    assert eng.blocks.num_free == 32


@pytest.mark.parametrize("prefix_cache", [False, True], ids=["swap-only", "swap+prefix"])
def test_tiny_swap_preemption_changes_nothing_cpu(prefix_cache):
    base, rb, _ = _tiny("cpu", OracleOps, 64, None, max_new=24)
    # 5 requests need ~20 blocks at their longest; 12 force preemption during decode
    eng, rs, mgr = _tiny("cpu", OracleOps, 12, dict(enable_prefix_cache=prefix_cache), max_new=24)
    _check_same(base, rb, eng, rs, 1e-4)
    assert mgr.stats.preemptions > 0
    assert mgr.stats.blocks_swapped_in == mgr.stats.blocks_swapped_out > 0
    assert not mgr.swapped
    assert eng.blocks.num_free == 12 and all(c == 0 for c in eng.blocks.ref_count)
    assert len(mgr.host_k.free_blocks) == 64                     # the swap space was given back


def test_install_needs_the_adapter_and_is_idempotent():
    eng = TinyEngine(device="cpu", hashed_blocks=True)
    with pytest.raises(RuntimeError, match="engine_adapter.install"):
        kv_manager.install(eng, ops=OracleOps)
    engine_adapter.install(eng, ops=OracleOps)
    m = kv_manager.install(eng, ops=OracleOps)
    assert kv_manager.install(eng, ops=OracleOps) is m
    kv_manager.uninstall(eng)
    assert eng._pa_b200_kv is None


# ---------------------------------------------------------------------------------------
# the real reference engine (CPU, build container only)
# ---------------------------------------------------------------------------------------
def _ref_engine(mv, state, num_kv_blocks, manager, budget=32):
    torch.manual_seed(0)
    eng = mv.MiniVLLM(vocab_size=150, hidden_size=64, num_layers=2, num_heads=4, intermediate_size=128,
                      num_kv_blocks=num_kv_blocks, block_size=4, max_num_seqs=6, max_num_batched_tokens=budget)
    if state is not None:
        weights = {k: v.clone() for k, v in state.items() if k not in ("k_cache", "v_cache")}
        eng.model.load_state_dict(weights, strict=False)            # the KV buffers are sized per engine
    engine_adapter.install(eng, ops=OracleOps)
    mgr = kv_manager.install(eng, num_host_blocks=128, ops=OracleOps, **manager) if manager is not None else None
    return eng, mgr


def _ref_run(eng, prompts, max_new):
    """Greedy generation, one new request per step after the first two steps; returns the tokens
    of every request in submission order."""
    outs, pending = {}, list(prompts)
    rid0 = eng.add_request(pending.pop(0), max_new_tokens=max_new, temperature=0.0)
    order = [rid0]
    steps = 0
    while (eng.scheduler.has_unfinished or pending) and steps < 600:
        if pending and steps >= 2:
            order.append(eng.add_request(pending.pop(0), max_new_tokens=max_new, temperature=0.0))
        res = eng.step()
        for rid, tok in res:
            outs.setdefault(rid, []).append(tok)
        steps += 1
    assert not eng.scheduler.has_unfinished
    return [outs[r] for r in order]


REF_PREFIX = [9, 8, 7, 6, 5, 4, 3, 2, 11, 12, 13]                # 2 full blocks of 4 + 3 tokens
REF_PROMPTS = [REF_PREFIX + [20 + i, 40 + i] * (3 + i) for i in range(4)] + [list(range(60, 83))]


@needs_reference
def test_reference_engine_prefix_cache_changes_nothing():
    mv = _ref_module()
    base, _ = _ref_engine(mv, None, 128, None)
    state = base.model.state_dict()
    want = _ref_run(base, REF_PROMPTS, 8)
    eng, mgr = _ref_engine(mv, state, 128, dict(enable_swap=False))
    got = _ref_run(eng, REF_PROMPTS, 8)
    assert got == want
    assert mgr.stats.prefix_hit_blocks == 3 * 2                   # requests 1..3 share REF_PREFIX's 2 full blocks
    assert eng.scheduler.allocator.num_free == 128
    assert all(c == 0 for c in eng.scheduler.allocator.ref_count)


@needs_reference
def test_reference_engine_swap_preemption_changes_nothing():
    mv = _ref_module()
    base, _ = _ref_engine(mv, None, 128, None)
    state = base.model.state_dict()
    want = _ref_run(base, REF_PROMPTS, 16)
    # the five requests need ~45 blocks of 4 at their longest; 24 force preemption
    eng, mgr = _ref_engine(mv, state, 24, dict())
    got = _ref_run(eng, REF_PROMPTS, 16)
    assert got == want
    assert mgr.stats.preemptions > 0 and mgr.stats.blocks_swapped_in == mgr.stats.blocks_swapped_out
    assert not mgr.swapped and eng.scheduler.allocator.num_free == 24


@needs_reference
def test_reference_engine_without_manager_stalls_where_manager_swaps():
    """The gap being closed: with a full pool the reference scheduler skips every request
    (mini_vllm.py:231-233) and ``generate`` never ends; the manager preempts instead."""
    mv = _ref_module()
    eng, _ = _ref_engine(mv, None, 24, None)
    for p in REF_PROMPTS:
        eng.add_request(p, max_new_tokens=16, temperature=0.0)
    stuck = 0
    for _ in range(300):
        if not eng.scheduler.has_unfinished:
            break
        stuck = stuck + 1 if not eng.step() else 0
        if stuck >= 5:
            break
    assert stuck >= 5, "expected the unmanaged engine to stall on a full pool"


# ---------------------------------------------------------------------------------------
# GPU: the same scenarios through the CUDA kernels (swap = pa_swap_blocks*, attention = pa_decode /
# pa_prefill over shared and relocated blocks)
# ---------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("cfg", [dict(hidden=128, heads=4), dict(hidden=256, heads=2)], ids=["d32", "d128"])
@pytest.mark.parametrize("manager,blocks", [(dict(enable_swap=False), 64), (dict(enable_prefix_cache=False), 12), (dict(), 12)],
                         ids=["prefix", "swap", "prefix+swap"])
def test_tiny_engine_cuda_managed_matches_unmanaged_oracle_engine(cfg, manager, blocks):
    from vllm_from_scratch_b200 import paged_attention as P
    cpu, rc, _ = _tiny("cpu", OracleOps, 64, None, max_new=24, **cfg)
    gpu, rg, mgr = _tiny("cuda", P, blocks, manager, max_new=24, **cfg)
    for x, y in zip(rc, rg):
        assert x.output_token_ids == y.output_token_ids, (x.request_id, x.output_token_ids, y.output_token_ids)
        for la, lb in zip(cpu.logits_log[x.request_id], gpu.logits_log[y.request_id]):
            assert torch.allclose(la, lb, atol=2e-4, rtol=1e-4), (la - lb).abs().max()
    if manager.get("enable_prefix_cache", True):
        assert mgr.stats.prefix_hit_blocks >= 2
    if manager.get("enable_swap", True):
        assert mgr.stats.preemptions > 0 and mgr.stats.blocks_swapped_in == mgr.stats.blocks_swapped_out
        assert not mgr.swapped and len(mgr.host_k.free_blocks) == 64
    assert gpu.blocks.num_free == blocks and all(c == 0 for c in gpu.blocks.ref_count)


@pytest.mark.gpu
def test_tiny_engine_cuda_bf16_swap_through_tensor_core_path():
    """bf16, head_dim 128, block_size 16: every step is one tcgen05 varlen launch; preempted
    requests come back into other blocks and keep generating what the unmanaged engine generates
    (first-token logits within the bf16 tolerance; greedy tokens may flip on ties later on)."""
    from vllm_from_scratch_b200 import paged_attention as P
    cfg = dict(dtype=torch.bfloat16, hidden=256, heads=2, layers=1)
    a, ra, _ = _tiny("cuda", P, 64, None, max_new=24, **cfg)
    b, rb, mgr = _tiny("cuda", P, 12, dict(enable_prefix_cache=False), max_new=24, **cfg)
    assert mgr.stats.preemptions > 0 and not mgr.swapped
    same = 0
    for x, y in zip(ra, rb):
        la, lb = a.logits_log[x.request_id][0], b.logits_log[y.request_id][0]
        assert torch.allclose(la, lb, atol=5e-2), (la - lb).abs().max()
        same += x.output_token_ids == y.output_token_ids
    assert same >= len(ra) // 2, (same, len(ra))
    assert b.blocks.num_free == 12
