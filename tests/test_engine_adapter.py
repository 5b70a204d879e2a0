"""Engine call site (config C4): the adapter under the engine loop.

CPU part (build container): the adapter's plumbing against the REAL reference engine
(/root/reference/05_mini_vllm/mini_vllm.py, imported unmodified; skipped where the reference is
absent) with the oracle as kernel namespace, and against the TinyEngine harness.
GPU part: TinyEngine on cuda through the CUDA kernels vs the same engine on CPU through the
oracle - mixed chunked-prefill + decode batches, shared prefix blocks, swap-out / swap-in and
block copy in the middle of generation.
"""
import copy
import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from engine_harness import OracleOps, TinyEngine  # noqa: E402
from vllm_from_scratch_b200 import engine_adapter  # noqa: E402

REF_DIR = "/root/reference/05_mini_vllm"
needs_reference = pytest.mark.skipif(not os.path.exists(os.path.join(REF_DIR, "mini_vllm.py")),
                                     reason="the reference tree is only present in the build container")


def _ref_module():
    if REF_DIR not in sys.path:
        sys.path.insert(0, REF_DIR)
    import mini_vllm
    return mini_vllm


def _capture_logits(engine):
    """Record the logits of every model call (in call order)."""
    log = []
    orig = engine.model.forward

    def fwd(*a, **k):
        out = orig(*a, **k)
        log.append(out.detach().float().cpu().clone())
        return out

    engine.model.forward = fwd
    return log


# ---------------------------------------------------------------------------------------
# real reference engine, CPU, oracle-backed ops
# ---------------------------------------------------------------------------------------
@needs_reference
def test_reference_engine_runs_with_adapter_and_restores():
    mv = _ref_module()
    torch.manual_seed(0)
    eng = mv.MiniVLLM(vocab_size=100, hidden_size=64, num_layers=2, num_heads=4, intermediate_size=128,
                      num_kv_blocks=64, block_size=4, max_num_seqs=4, max_num_batched_tokens=16)
    orig_forward = eng.model.layers[0].attn.forward
    engine_adapter.install(eng, ops=OracleOps)
    assert engine_adapter.install(eng, ops=OracleOps) is eng            # idempotent
    outs = eng.generate([[5, 6, 7, 8, 9, 10, 11], [3, 4], list(range(1, 30))], max_new_tokens=6)
    assert len(outs) == 3 and all(1 <= len(v) <= 6 for v in outs.values())
    assert eng.allocator.num_free_blocks == 64 if hasattr(eng.allocator, "num_free_blocks") else True
    engine_adapter.uninstall(eng)
    assert eng.model.layers[0].attn.forward == orig_forward
    assert engine_adapter.current_metadata(eng) is None


@needs_reference
def test_unchunked_prefill_matches_reference_own_prefill_branch():
    """For one un-chunked prompt the reference's prefill branch is exact causal attention
    (mini_vllm.py:437-458): the adapter must reproduce its logits."""
    mv = _ref_module()
    torch.manual_seed(1)
    kw = dict(vocab_size=120, hidden_size=64, num_layers=2, num_heads=4, intermediate_size=128,
              num_kv_blocks=32, block_size=4, max_num_seqs=2, max_num_batched_tokens=256)
    ref = mv.MiniVLLM(**kw)
    ours = mv.MiniVLLM(**kw)
    ours.model.load_state_dict(copy.deepcopy(ref.model.state_dict()))
    engine_adapter.install(ours, ops=OracleOps)
    prompt = [7, 3, 99, 15, 4, 4, 28, 51, 60, 2, 13]
    la, lb = _capture_logits(ref), _capture_logits(ours)
    ref.add_request(prompt, max_new_tokens=1)
    ours.add_request(prompt, max_new_tokens=1)
    ref.step()
    ours.step()
    assert torch.allclose(la[0], lb[0], atol=1e-5), (la[0] - lb[0]).abs().max()


@needs_reference
def test_stepwise_decode_equals_full_causal_forward():
    """The property SURVEY.md 8(c) names: with the adapter, the logits of decode step n equal
    the last-token logits of one causal forward over prompt + the n generated tokens."""
    mv = _ref_module()
    torch.manual_seed(2)
    kw = dict(vocab_size=150, hidden_size=64, num_layers=2, num_heads=4, intermediate_size=128,
              num_kv_blocks=64, block_size=4, max_num_seqs=2, max_num_batched_tokens=256)
    eng = mv.MiniVLLM(**kw)
    engine_adapter.install(eng, ops=OracleOps)
    log = _capture_logits(eng)
    prompt = [9, 8, 7, 6, 5, 4, 3, 2, 1, 11, 12, 13, 14]
    rid = eng.add_request(prompt, max_new_tokens=5, temperature=0.0) if "temperature" in mv.MiniVLLM.add_request.__code__.co_varnames \
        else eng.add_request(prompt, max_new_tokens=5)
    gen = []
    while eng.scheduler.has_unfinished:
        for r, tok in eng.step():
            if r == rid:
                gen.append(tok)
    assert len(gen) >= 2
    # one causal forward over prompt + gen[:-1] through the reference's own (exact) prefill branch
    full = mv.MiniVLLM(**kw)
    full.model.load_state_dict(copy.deepcopy(eng.model.state_dict()))
    lf = _capture_logits(full)
    full.add_request(prompt + gen[:len(log) - 1], max_new_tokens=1)
    full.step()
    assert torch.allclose(log[-1][0], lf[0][0], atol=2e-5), (log[-1][0] - lf[0][0]).abs().max()


# ---------------------------------------------------------------------------------------
# TinyEngine harness
# ---------------------------------------------------------------------------------------
PROMPTS = [list(range(3, 3 + n)) for n in (70, 9, 33, 48, 17)]


def _run_tiny(device, ops, dtype=torch.float32, token_budget=48, share=True, mid_hook=None, seed=0, **kw):
    eng = TinyEngine(device=device, dtype=dtype, token_budget=token_budget, seed=seed, **kw)
    engine_adapter.install(eng, ops=ops)
    reqs = [eng.add_request(p, max_new_tokens=6) for p in PROMPTS]
    steps = 0
    shared_added = False
    while eng.scheduler.has_unfinished and steps < 200:
        eng.step()
        steps += 1
        if share and not shared_added and reqs[0].num_computed_tokens >= 32:
            # a request whose first 32 tokens equal request 0's: prefix-cache hit on 2 blocks
            r = eng.add_request(PROMPTS[0][:32] + [150, 151, 152, 153, 154], max_new_tokens=6,
                                shared_prefix_of=reqs[0], shared_tokens=32)
            reqs.append(r)
            shared_added = True
        if mid_hook is not None:
            mid_hook(eng, steps)
    assert not eng.scheduler.has_unfinished
    return eng, reqs


def test_tiny_engine_chunked_and_shared_equals_unchunked_cpu():
    a, ra = _run_tiny("cpu", OracleOps, token_budget=48, share=False)
    b, rb = _run_tiny("cpu", OracleOps, token_budget=4096, share=False)
    for x, y in zip(ra, rb):
        assert x.output_token_ids == y.output_token_ids
        for la, lb in zip(a.logits_log[x.request_id], b.logits_log[y.request_id]):
            assert torch.allclose(la, lb, atol=1e-4)
    # shared prefix: the late request produces what it would produce alone
    c, rc = _run_tiny("cpu", OracleOps, token_budget=48, share=True)
    solo = TinyEngine(device="cpu", token_budget=4096)
    engine_adapter.install(solo, ops=OracleOps)
    r = solo.add_request(PROMPTS[0][:32] + [150, 151, 152, 153, 154], max_new_tokens=6)
    solo.run()
    assert rc[-1].output_token_ids == r.output_token_ids
    assert all(rc[i].output_token_ids == ra[i].output_token_ids for i in range(len(ra)))
    assert sorted(c.blocks.free) == list(range(64))          # every block returned exactly once


@pytest.mark.gpu
@pytest.mark.parametrize("cfg", [dict(hidden=128, heads=4), dict(hidden=256, heads=2, block_size=16)],
                         ids=["d32_generic", "d128_generic_f32"])
def test_tiny_engine_cuda_matches_oracle_engine(cfg):
    """C4: mixed chunked-prefill + decode steps with prefix sharing, swap-out / swap-in and a block
    copy in the middle of generation - CUDA kernels vs the CPU oracle, same weights."""
    from vllm_from_scratch_b200 import paged_attention as P
    state = {}

    def swap_and_copy(eng, step):
        if eng.device.type != "cuda":
            return
        if step == 3:
            # swap request 1 out to pinned host memory, scribble over its device blocks, swap it back
            # into DIFFERENT blocks, and patch its block table - generation must not notice
            r = [q for q in eng.scheduler.running if q.request_id == "req-1"]
            if not r:
                return
            r = r[0]
            caches = [eng.model.k_cache, eng.model.v_cache]
            host = [P.HostKVPool(c, num_blocks=8, block_dim=1) for c in caches]
            pairs = [(b, i) for i, b in enumerate(r.block_table)]
            evs = [P.swap_out(c, h.kv_cache, pairs, block_dim=1) for c, h in zip(caches, host)]
            for ev in evs:
                ev.synchronize()
            for c in caches:
                c[:, r.block_table] = float("nan")
            new = [eng.blocks.take() for _ in r.block_table]
            back = [(i, b) for i, b in enumerate(new)]
            for c, h in zip(caches, host):
                torch.cuda.current_stream().wait_event(P.swap_in(h.kv_cache, c, back, block_dim=1))
            for b in r.block_table:
                eng.blocks.release(b)
            r.block_table[:] = new
            state["swapped"] = True
        if step == 5:
            # copy-on-write style block copy: move request 2's first block elsewhere
            r = [q for q in eng.scheduler.running if q.request_id == "req-2"]
            if not r:
                return
            r = r[0]
            dst = eng.blocks.take()
            P.copy_blocks([eng.model.k_cache, eng.model.v_cache], [(r.block_table[0], dst)], block_dim=1)
            eng.blocks.release(r.block_table[0])
            r.block_table[0] = dst
            state["copied"] = True

    cpu, rc = _run_tiny("cpu", OracleOps, **cfg)
    gpu, rg = _run_tiny("cuda", None, mid_hook=swap_and_copy, **cfg)
    assert state.get("swapped") and state.get("copied")
    assert len(rc) == len(rg) == len(PROMPTS) + 1
    for x, y in zip(rc, rg):
        assert x.output_token_ids == y.output_token_ids, (x.request_id, x.output_token_ids, y.output_token_ids)
        for la, lb in zip(cpu.logits_log[x.request_id], gpu.logits_log[y.request_id]):
            assert torch.allclose(la, lb, atol=2e-4, rtol=1e-4), (la - lb).abs().max()


@pytest.mark.gpu
def test_tiny_engine_cuda_bf16_fast_path():
    """bf16 engine with head_dim 128: the decode steps go through the TMA fast path on the
    engine's 'bshd' layout.  Checked at the op level inside the loop: every attention call's output
    vs the oracle on the same inputs (bf16 tolerance 2e-2)."""
    from vllm_from_scratch_b200 import paged_attention as P
    worst = {"decode": 0.0, "prefill": 0.0, "calls": 0}

    class Checked:
        reshape_and_cache = staticmethod(P.reshape_and_cache)
        block_table_step = staticmethod(P.block_table_step)

        @staticmethod
        def paged_attention_decode_batched(q, cache, bts, lens):
            out = P.paged_attention_decode_batched(q, cache, bts, lens)
            ref = OracleOps.paged_attention_decode_batched(q.float().cpu(), tuple(c.float().cpu() for c in cache),
                                                           bts.cpu(), lens.cpu())
            worst["decode"] = max(worst["decode"], (out.float().cpu() - ref).abs().max().item())
            worst["calls"] += 1
            return out

        @staticmethod
        def paged_attention_prefill_chunk(q, cache, bt, start):
            out = P.paged_attention_prefill_chunk(q, cache, bt, start)
            ref = OracleOps.paged_attention_prefill_chunk(q.float().cpu(), tuple(c.float().cpu() for c in cache),
                                                          bt.cpu(), start)
            worst["prefill"] = max(worst["prefill"], (out.float().cpu() - ref).abs().max().item())
            return out

    eng, reqs = _run_tiny("cuda", Checked, dtype=torch.bfloat16, hidden=256, heads=2, layers=1)
    assert worst["calls"] > 0
    assert worst["decode"] <= 2e-2 and worst["prefill"] <= 2e-2, worst


@pytest.mark.gpu
def test_tiny_engine_cuda_bf16_varlen_one_launch():
    """bf16, head_dim 128, block_size 16: the adapter sends every mixed chunked-prefill + decode
    step through ONE pa_prefill_varlen launch (tcgen05 kernel).  Each call is checked against the
    oracle per request, and the generated tokens must equal those of the two-kernel route."""
    from vllm_from_scratch_b200 import paged_attention as P
    worst = {"err": 0.0, "calls": 0, "mixed": 0}

    class Varlen:
        reshape_and_cache = staticmethod(P.reshape_and_cache)
        block_table_step = staticmethod(P.block_table_step)
        prefill_tensor_core_available = staticmethod(P.prefill_tensor_core_available)

        @staticmethod
        def paged_attention_prefill_varlen(q, cache, bts, cu, ctx, max_q):
            out = P.paged_attention_prefill_varlen(q, cache, bts, cu, ctx, max_q)
            cu_h, ctx_h = cu.tolist(), ctx.tolist()
            cpu_cache = tuple(c.float().cpu() for c in cache)
            lens = [cu_h[i + 1] - cu_h[i] for i in range(len(ctx_h))]
            worst["mixed"] += int(min(lens) == 1 and max(lens) > 1)
            for i, n in enumerate(lens):
                ref = OracleOps.paged_attention_prefill_chunk(q[cu_h[i]:cu_h[i + 1]].float().cpu(), cpu_cache,
                                                              bts[i].cpu(), ctx_h[i] - n)
                worst["err"] = max(worst["err"], (out[cu_h[i]:cu_h[i + 1]].float().cpu() - ref).abs().max().item())
            worst["calls"] += 1
            return out

    class TwoKernel:
        reshape_and_cache = staticmethod(P.reshape_and_cache)
        block_table_step = staticmethod(P.block_table_step)
        paged_attention_decode_batched = staticmethod(P.paged_attention_decode_batched)
        paged_attention_prefill_chunk = staticmethod(P.paged_attention_prefill_chunk)

    cfg = dict(dtype=torch.bfloat16, hidden=256, heads=2, layers=1, block_size=16)
    a, ra = _run_tiny("cuda", Varlen, **cfg)
    assert worst["calls"] > 0 and worst["mixed"] > 0, worst
    assert worst["err"] <= 2e-2, worst
    b, rb = _run_tiny("cuda", TwoKernel, **cfg)
    same = sum(x.output_token_ids == y.output_token_ids for x, y in zip(ra, rb))
    assert same >= len(ra) // 2, (same, len(ra))    # greedy tokens of a random-init bf16 model may flip on ties


@pytest.mark.gpu
def test_tiny_engine_cuda_bf16_split_route_for_decode_heavy_mixed_steps(monkeypatch):
    """Mixed steps whose decode requests stream a lot of KV go: decode rows -> decode kernel, prefill
    chunks -> the same varlen launch with the decode requests masked out (ctx 0).  Forced here with a
    zero threshold; every row of the step must match the one-launch route within the bf16 tolerance."""
    from vllm_from_scratch_b200 import paged_attention as P
    seen = {"split": 0, "err": 0.0}

    class Checked:
        reshape_and_cache = staticmethod(P.reshape_and_cache)
        block_table_step = staticmethod(P.block_table_step)
        prefill_tensor_core_available = staticmethod(P.prefill_tensor_core_available)
        paged_attention_decode_batched = staticmethod(P.paged_attention_decode_batched)

        @staticmethod
        def paged_attention_prefill_varlen(q, cache, bts, cu, ctx, max_q, out=None):
            res = P.paged_attention_prefill_varlen(q, cache, bts, cu, ctx, max_q, out=out)
            if out is not None:                              # split route: decode requests are masked out
                seen["split"] += 1
                assert int((ctx == 0).sum()) > 0
                Checked.pending = (q, cache, bts, cu, max_q, res)
            return res

    orig_attend = engine_adapter._attend

    def attend(attn, ops, state, md, q, cache, t, h, d):
        y = orig_attend(attn, ops, state, md, q, cache, t, h, d)
        if getattr(Checked, "pending", None) is not None:
            qq, cc, bts, cu, max_q, out = Checked.pending
            Checked.pending = None
            full = P.paged_attention_prefill_varlen(qq, cc, bts, cu, md.ctx_lens, max_q)     # one-launch route
            seen["err"] = max(seen["err"], (full.float() - out.float()).abs().max().item())
        return y

    monkeypatch.setattr(engine_adapter, "_attend", attend)
    monkeypatch.setattr(engine_adapter, "SPLIT_DECODE_BYTES", 0)
    Checked.pending = None
    _run_tiny("cuda", Checked, dtype=torch.bfloat16, hidden=256, heads=2, layers=1, block_size=16)
    assert seen["split"] > 0, "no mixed step took the split route"
    assert seen["err"] <= 2e-2, seen
