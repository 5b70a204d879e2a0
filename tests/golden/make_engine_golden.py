"""Record the REAL reference engine loop (config C4) as a golden step trace.

Usage (build container only - /root/reference does not exist on the GPU box):

    python tests/golden/make_engine_golden.py

For every case of ``engine_cases.ENGINE_CASES`` this imports the UNMODIFIED
``/root/reference/05_mini_vllm/mini_vllm.py`` (read-only, through ``sys.path``), builds
``MiniVLLM`` (:641-691) on the CPU, loads the seeded weights of ``engine_cases.det_weights``,
installs ``engine_adapter`` + ``kv_manager`` with the CPU ORACLE as kernel namespace, feeds the
seeded prompts (greedy sampling) with ``add_request`` (:693-710) and drives ``step()`` (:712-764)
until every request has finished.  Recorded per step, exactly as the real ``Scheduler.schedule``
(:221-310) returned them under the real ``BlockAllocator`` (:32-108): the scheduled requests' tokens,
start positions, slot mappings, prefill flags and block tables; the swap-out / swap-in block pairs
``kv_manager`` issued before the step; and the logits the real ``MiniTransformer.forward``
(:554-585) produced (top-8 of every row + greedy token for all steps, full rows for every third step).

``tests/test_engine_golden.py`` replays the trace on the B200 through the CUDA kernels
(``tests/engine_replica.py``) and compares the logits step by step.
"""
import hashlib
import os
import sys

import numpy as np
import torch

REF = os.environ.get("PA_REFERENCE_ROOT", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "tests"), HERE):
    if p not in sys.path:
        sys.path.insert(0, p)

from engine_cases import ENGINE_CASES, det_weights, engine_kwargs, make_prompts, torch_dtype  # noqa: E402
from engine_harness import OracleOps  # noqa: E402
from vllm_from_scratch_b200 import engine_adapter, kv_manager  # noqa: E402


def weights_sha(weights) -> str:
    h = hashlib.sha256()
    for name in sorted(weights):
        h.update(name.encode())
        h.update(weights[name].contiguous().numpy().tobytes())
    return h.hexdigest()


def record(name, cfg, mv):
    torch.manual_seed(cfg["seed"])
    eng = mv.MiniVLLM(**engine_kwargs(cfg))
    weights = det_weights(cfg)
    res = eng.model.load_state_dict(weights, strict=False)
    assert not res.unexpected_keys, res
    assert all(k.endswith(("_cache", "_cached", "inv_freq")) for k in res.missing_keys), res
    eng.model.to(torch_dtype(cfg))

    events = []

    class LoggingOps(OracleOps):
        """Oracle ops that log the swap traffic and poison swapped-out device blocks."""

        @staticmethod
        def swap_out(dev_cache, host_cache, pairs, block_dim=None):
            OracleOps.swap_out(dev_cache, host_cache, pairs, block_dim)
            # poison what nobody may read again: blocks registered under a prefix hash stay valid cache
            # candidates after the free (mini_vllm.py:92-96) and keep their content
            dead = [b for b, _ in pairs if b not in eng.allocator.block_to_hash]
            events.append((0, 0 if dev_cache is eng.model.k_cache else 1, [tuple(p) for p in pairs], dead))
            if dead:
                dev_cache[:, dead] = float("nan")

        @staticmethod
        def swap_in(host_cache, dev_cache, pairs, block_dim=None):
            OracleOps.swap_in(host_cache, dev_cache, pairs, block_dim)
            events.append((1, 0 if dev_cache is eng.model.k_cache else 1, [tuple(p) for p in pairs], []))

    engine_adapter.install(eng, ops=LoggingOps)
    mgr = kv_manager.install(eng, num_host_blocks=cfg["num_host_blocks"], ops=LoggingOps)

    steps, logits_log = [], []
    inner = eng.scheduler.schedule
    seen_events = [0]

    def recording_schedule():
        out = inner()
        if out:
            steps.append({"events": events[seen_events[0]:],
                          "sched": [(int(s.request.request_id.split("-")[1]), list(s.tokens), int(s.positions[0]),
                                     bool(s.is_prefill), list(s.slot_mapping), list(s.request.block_table))
                                    for s in out]})
            seen_events[0] = len(events)
        return out

    eng.scheduler.schedule = recording_schedule
    fwd = eng.model.forward

    def recording_forward(*a, **k):
        out = fwd(*a, **k)
        logits_log.append(out.detach().float().clone())
        return out

    eng.model.forward = recording_forward

    prompts, arrival = make_prompts(cfg)
    pending = sorted(range(len(prompts)), key=lambda i: arrival[i])
    outputs = {}
    n = 0
    while (eng.scheduler.has_unfinished or pending) and n < 2000:
        while pending and arrival[pending[0]] <= n:
            i = pending.pop(0)
            rid = eng.add_request(prompts[i], max_new_tokens=cfg["max_new_tokens"], temperature=0.0)
            assert rid == f"req-{i}", (rid, i)
        for rid, tok in eng.step():
            outputs.setdefault(rid, []).append(tok)
        n += 1
    assert not eng.scheduler.has_unfinished and not pending, "the engine did not drain"
    assert len(steps) == len(logits_log)
    assert all(torch.isfinite(l).all() for l in logits_log), "a poisoned (swapped-out) block was read"

    # ---- flatten -------------------------------------------------------------------------
    step_ptr, req_ids, start_pos, is_prefill = [0], [], [], []
    tok_ptr, tokens_flat, slots_flat, bt_ptr, bt_flat = [0], [], [], [0], []
    ev_step, ev_kind, ev_cache, ev_ptr, ev_pairs = [], [], [], [0], []
    ev_dead_ptr, ev_dead = [0], []
    next_tok, top_val, top_idx = [], [], []
    full_steps, full_ptr, full_logits = [], [0], []
    for s, (st, lg) in enumerate(zip(steps, logits_log)):
        for kind, cache, pairs, dead in st["events"]:
            ev_step.append(s)
            ev_kind.append(kind)
            ev_cache.append(cache)
            ev_pairs.extend(pairs)
            ev_ptr.append(len(ev_pairs))
            ev_dead.extend(dead)
            ev_dead_ptr.append(len(ev_dead))
        assert lg.shape[0] == len(st["sched"])
        for (rid, toks, start, pf, slots, bt), row in zip(st["sched"], lg):
            req_ids.append(rid)
            start_pos.append(start)
            is_prefill.append(pf)
            tokens_flat.extend(toks)
            slots_flat.extend(slots)
            tok_ptr.append(len(tokens_flat))
            bt_flat.extend(bt)
            bt_ptr.append(len(bt_flat))
            tv, ti = row.topk(8)
            next_tok.append(int(row.argmax()))
            top_val.append(tv.numpy())
            top_idx.append(ti.numpy())
        step_ptr.append(len(req_ids))
        if s % 3 == 0:
            full_steps.append(s)
            full_logits.append(lg.numpy())
            full_ptr.append(full_ptr[-1] + lg.shape[0])
    st = mgr.stats
    out = dict(
        step_ptr=np.asarray(step_ptr, np.int32), req_ids=np.asarray(req_ids, np.int32),
        start_pos=np.asarray(start_pos, np.int32), is_prefill=np.asarray(is_prefill, np.bool_),
        tok_ptr=np.asarray(tok_ptr, np.int32), tokens_flat=np.asarray(tokens_flat, np.int32),
        slots_flat=np.asarray(slots_flat, np.int32), bt_ptr=np.asarray(bt_ptr, np.int32),
        bt_flat=np.asarray(bt_flat, np.int32),
        ev_step=np.asarray(ev_step, np.int32), ev_kind=np.asarray(ev_kind, np.int8),
        ev_cache=np.asarray(ev_cache, np.int8), ev_ptr=np.asarray(ev_ptr, np.int32),
        ev_pairs=np.asarray(ev_pairs, np.int32).reshape(-1, 2),
        ev_dead_ptr=np.asarray(ev_dead_ptr, np.int32), ev_dead=np.asarray(ev_dead, np.int32),
        next_tok=np.asarray(next_tok, np.int32), top_val=np.stack(top_val).astype(np.float32),
        top_idx=np.stack(top_idx).astype(np.int32),
        full_steps=np.asarray(full_steps, np.int32), full_ptr=np.asarray(full_ptr, np.int32),
        full_logits=np.concatenate(full_logits).astype(np.float32),
        weights_sha=np.asarray(weights_sha(weights)),
        stats=np.asarray([st.prefix_hit_blocks, st.prefix_hit_tokens, st.preemptions, st.blocks_swapped_out,
                          st.blocks_swapped_in], np.int32),
        outputs=np.asarray([len(outputs.get(f"req-{i}", [])) for i in range(len(prompts))], np.int32),
    )
    path = os.path.join(HERE, f"engine_{name}.npz")
    np.savez_compressed(path, **out)
    mixed = sum(1 for s in steps if any(x[3] for x in s["sched"]) and any(not x[3] for x in s["sched"]))
    print(f"{name}: {len(steps)} steps ({mixed} mixed prefill + decode), {len(req_ids)} scheduled requests, "
          f"{len(tokens_flat)} tokens, prefix hits {st.prefix_hit_blocks} blocks, preemptions {st.preemptions} "
          f"({st.blocks_swapped_out} blocks out / {st.blocks_swapped_in} in), {os.path.getsize(path)} bytes")
    assert st.prefix_hit_blocks > 0 and st.preemptions > 0 and mixed > 0, "the trace must exercise sharing, swap and mixed steps"


if __name__ == "__main__":
    sys.path.insert(0, os.path.join(REF, "05_mini_vllm"))
    import mini_vllm as mv  # noqa: E402  (the unmodified reference module)
    only = sys.argv[1:] or list(ENGINE_CASES)
    for name in only:
        record(name, ENGINE_CASES[name], mv)
