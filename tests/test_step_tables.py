"""Row f3: persistent device block tables (``engine_adapter.StepTables`` + ``pa_block_table_step``) and the
CUDA-graph replay of decode-only steps (``engine_adapter.DecodeGraphs``).

The expected side of every table check is the reference engine's own per-step rebuild
(05_mini_vllm/mini_vllm.py:729-739: the step's tensors are rebuilt from each request's ``block_table``
list): after any history of appends, replacements (swap-in), evictions and capacity growth, the gathered
``[B, W]`` table must hold exactly ``request.block_table[:ceil(ctx / BS)]`` per scheduled request - bit-exact.
CPU tests run the adapter logic over the oracle's ``block_table_step``; GPU tests run the CUDA kernel.
"""
import os
import random
import sys
from dataclasses import dataclass, field
from typing import List

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from engine_harness import OracleOps, TinyEngine  # noqa: E402
from oracle import paged_attention_oracle as O  # noqa: E402
from vllm_from_scratch_b200 import engine_adapter  # noqa: E402

BS = 4


@dataclass
class _Req:
    request_id: str
    block_table: List[int] = field(default_factory=list)
    ntok: int = 0


@dataclass
class _Sched:
    request: _Req
    tokens: List[int]
    positions: List[int]
    is_prefill: bool
    slot_mapping: List[int]


def _random_history(ops, device, steps=120, seed=0, max_seqs=4, max_blocks=4, max_tokens=8, token_tables=False):
    """Drive StepTables with a random scheduler: new requests, decode tokens, prefill chunks, block replacement
    (swap-in), requests that disappear for a while (row eviction) and more requests / longer tables than the
    initial capacity (growth).  After every step the device tables are compared with the lists."""
    rng = random.Random(seed)
    tables = engine_adapter.StepTables(device, ops, BS, max_seqs=max_seqs, max_tokens=max_tokens, max_blocks=max_blocks,
                                       token_tables=token_tables)
    live: List[_Req] = []
    next_block, nreq = 0, 0
    generations = set()
    for step in range(steps):
        if len(live) < 3 or rng.random() < 0.3:
            live.append(_Req(f"r{nreq}"))
            nreq += 1
        if len(live) > 12:
            live.pop(rng.randrange(len(live)))
        batch = rng.sample(live, rng.randint(1, min(len(live), 6)))
        scheduled = []
        for r in batch:
            prefill = r.ntok == 0 or rng.random() < 0.25
            n = rng.randint(1, 9) if prefill else 1
            if rng.random() < 0.08 and r.block_table:                  # swap-in: every block id replaced
                r.block_table[:] = [next_block + i for i in range(len(r.block_table))]
                next_block += len(r.block_table)
            if rng.random() < 0.05 and len(r.block_table) > 1:         # copy-on-write of one middle block
                r.block_table[rng.randrange(len(r.block_table))] = next_block
                next_block += 1
            start = r.ntok
            while len(r.block_table) * BS < start + n:
                r.block_table.append(next_block)
                next_block += 1
            pos = list(range(start, start + n))
            slots = [r.block_table[p // BS] * BS + p % BS for p in pos]
            scheduled.append(_Sched(r, [rng.randint(1, 99) for _ in pos], pos, prefill, slots))
            r.ntok += n
        md = tables.build(scheduled)
        generations.add(tables.generation)
        # ---- the gathered table equals the reference's rebuild from the lists --------------------------------
        got = md.all_block_tables.cpu()
        for i, s in enumerate(scheduled):
            need = (s.positions[-1] + 1 + BS - 1) // BS
            assert got[i, :need].tolist() == s.request.block_table[:need], (step, i)
        assert md.ctx_lens.tolist() == [s.positions[-1] + 1 for s in scheduled]
        cu = md.cu_q_lens.tolist()
        assert cu == [sum(len(s.tokens) for s in scheduled[:i]) for i in range(len(scheduled) + 1)]
        assert md.token_ids.tolist() == [t for s in scheduled for t in s.tokens]
        assert md.positions.tolist() == [p for s in scheduled for p in s.positions]
        assert md.slot_mapping.tolist() == [x for s in scheduled for x in s.slot_mapping]
        assert md.slot_mapping.tolist() == [x for s in scheduled
                                            for x in O.compute_slot_mapping_range(s.request.block_table, s.positions[0],
                                                                                  len(s.positions), BS)]
        assert md.last_idx.tolist() == [c - 1 for c in cu[1:]]
        dec = [i for i, s in enumerate(scheduled) if not s.is_prefill]
        if dec:
            assert md.decode_rows.tolist() == [cu[i] for i in dec]
            assert md.decode_ctx_lens.tolist() == [scheduled[i].positions[-1] + 1 for i in dec]
            dt = md.decode_block_tables.cpu()
            for j, i in enumerate(dec):
                need = (scheduled[i].positions[-1] + 1 + BS - 1) // BS
                assert dt[j, :need].tolist() == scheduled[i].request.block_table[:need]
        assert md.all_decode == (len(dec) == len(scheduled))
        if dec and len(dec) < len(scheduled):
            assert md.ctx_lens_prefill_only.tolist() == [0 if not s.is_prefill else s.positions[-1] + 1 for s in scheduled]
        for (off, n, start, row), s in zip(md.prefill_chunks, [s for s in scheduled if s.is_prefill]):
            assert (n, start) == (len(s.tokens), s.positions[0])
            assert row[:(start + n + BS - 1) // BS].tolist() == s.request.block_table[:(start + n + BS - 1) // BS]
        if token_tables:
            # per-token form: token t of a request is a decode "sequence" of length position + 1 over the request's table
            tt = md.tok_block_tables.cpu()
            want_lens = [p + 1 for s in scheduled for p in s.positions]
            assert md.tok_ctx_lens.tolist() == want_lens
            at = 0
            for s in scheduled:
                for p in s.positions:
                    need = (p + 1 + BS - 1) // BS
                    assert tt[at, :need].tolist() == s.request.block_table[:need]
                    at += 1
        # padding invariant the decode graphs rely on: beyond the batch rows are -1 / lengths 0 / slots -1
        b, t = len(scheduled), cu[-1]
        assert (tables.dv["row_of"][b:tables.cap].cpu() == -1).all()
        assert (tables.dv["gather_lens"][b:tables.cap].cpu() == 0).all()
        assert (tables.dv["slots"][t:].cpu() == -1).all()
    return tables, generations


def test_step_tables_random_history_cpu():
    tables, generations = _random_history(OracleOps, "cpu")
    assert len(generations) > 1, "the history never outgrew the initial capacity"
    assert tables.cap >= 6 and tables.W > 4


def test_step_tables_per_token_form_cpu():
    tables, generations = _random_history(OracleOps, "cpu", seed=5, token_tables=True)
    assert tables.token_tables and len(generations) > 1


def test_step_tables_ship_only_the_delta_cpu():
    """Decode steps append at most one block id per request every BS tokens: the number of table entries
    shipped over a long decode equals the number of blocks allocated, not steps x blocks."""
    tables = engine_adapter.StepTables("cpu", OracleOps, BS, max_seqs=4, max_tokens=16, max_blocks=64)
    reqs = [_Req(f"r{i}") for i in range(3)]
    nb = 0
    for step in range(100):
        scheduled = []
        for r in reqs:
            if len(r.block_table) * BS < r.ntok + 1:
                r.block_table.append(nb)
                nb += 1
            scheduled.append(_Sched(r, [1], [r.ntok], step == 0 and False, [r.block_table[r.ntok // BS] * BS + r.ntok % BS]))
            r.ntok += 1
        md = tables.build(scheduled)
        assert md.all_decode
    assert tables.updates_sent == nb == 3 * 25


def test_step_tables_row_eviction_and_return_cpu():
    """More distinct requests than rows over time: the least recently scheduled row is recycled, and a request
    that comes back after losing its row gets its whole table re-sent."""
    tables = engine_adapter.StepTables("cpu", OracleOps, BS, max_seqs=2, max_tokens=8, max_blocks=8)
    reqs = [_Req(f"r{i}", [10 * i + j for j in range(3)], 9) for i in range(9)]
    for rnd in range(3):
        for i in range(0, 9, 2):
            batch = reqs[i:i + 2]
            scheduled = [_Sched(r, [1], [r.ntok], False, [r.block_table[r.ntok // BS] * BS + r.ntok % BS]) for r in batch]
            md = tables.build(scheduled)
            for j, r in enumerate(batch):
                assert md.all_block_tables[j, :3].tolist() == r.block_table
    assert tables.rows == 4 and len(tables.row_of_key) <= 4


def test_scheduling_a_request_twice_raises_cpu():
    tables = engine_adapter.StepTables("cpu", OracleOps, BS, max_seqs=2, max_tokens=8, max_blocks=8)
    r = _Req("r0", [1], 0)
    s = _Sched(r, [1], [0], False, [4])
    with pytest.raises(ValueError):
        tables.build([s, s])


# ---------------------------------------------------------------------------------------
# GPU: the kernel, the adapter on the device, graph replay
# ---------------------------------------------------------------------------------------
@pytest.mark.gpu
def test_block_table_step_kernel_bit_exact():
    from vllm_from_scratch_b200 import paged_attention as P
    g = torch.Generator().manual_seed(0)
    for rows, width, bs, nupd, ngather in ((8, 16, 16, 40, 6), (64, 300, 16, 5000, 64), (3, 1, 4, 7, 3), (512, 2048, 16, 3000, 300)):
        table = torch.randint(0, 1 << 20, (rows, width), generator=g, dtype=torch.int32)
        cap = nupd + 17
        upd = torch.zeros(cap, 3, dtype=torch.int32)
        # unique (row, col) targets - the host never emits duplicates - plus out-of-range entries that must be skipped
        cells = torch.randperm(rows * width, generator=g)[:min(nupd, rows * width)]
        n = cells.numel()
        upd[:n, 0], upd[:n, 1] = (cells // width).int(), (cells % width).int()
        upd[:n, 2] = torch.randint(0, 1 << 20, (n,), generator=g, dtype=torch.int32)
        if n > 4:
            upd[1, 0], upd[2, 1], upd[3, 0] = rows, width, -1
        upd[n:, 2] = 12345                                                # beyond the count: never applied
        count = torch.tensor([n], dtype=torch.int32)
        row_of = torch.randint(-1, rows, (ngather,), generator=g, dtype=torch.int32)
        lens = torch.randint(0, width * bs + 40, (ngather,), generator=g, dtype=torch.int32)
        step_ref = torch.full((ngather, width + 3), -7, dtype=torch.int32)
        table_ref = table.clone()
        O.block_table_step(table_ref, upd, count, row_of, lens, step_ref, bs)
        dev = "cuda"
        table_d, step_d = table.to(dev), torch.full((ngather, width + 3), -7, dtype=torch.int32, device=dev)
        P.block_table_step(table_d, upd.to(dev), count.to(dev), row_of.to(dev), lens.to(dev), step_d, bs)
        assert torch.equal(table_d.cpu(), table_ref)
        assert torch.equal(step_d.cpu(), step_ref)
    # update-only and gather-only calls
    t = torch.zeros(4, 4, dtype=torch.int32, device="cuda")
    P.block_table_step(t, torch.tensor([[1, 2, 9]], dtype=torch.int32, device="cuda"),
                       torch.tensor([1], dtype=torch.int32, device="cuda"), None, None, None, 16)
    assert t[1, 2].item() == 9 and int(t.sum()) == 9
    st = torch.zeros(2, 4, dtype=torch.int32, device="cuda")
    P.block_table_step(t, None, None, torch.tensor([1, -1], dtype=torch.int32, device="cuda"),
                       torch.tensor([64, 64], dtype=torch.int32, device="cuda"), st, 16)
    assert st.tolist() == [[0, 0, 9, 0], [0, 0, 0, 0]]


@pytest.mark.gpu
def test_step_tables_random_history_cuda():
    from vllm_from_scratch_b200 import paged_attention as P
    tables, generations = _random_history(P, "cuda", steps=150, seed=3)
    assert len(generations) > 1
    tables, generations = _random_history(P, "cuda", steps=100, seed=4, token_tables=True)
    assert tables.token_tables and len(generations) > 1


def _run(device, ops, graphs, prompts, dtype=torch.float32, new_tokens=12, **kw):
    eng = TinyEngine(device=device, dtype=dtype, **kw)
    engine_adapter.install(eng, ops=ops, cuda_graphs=graphs)
    reqs = [eng.add_request(p, max_new_tokens=new_tokens + i) for i, p in enumerate(prompts)]
    eng.run()
    assert not eng.scheduler.has_unfinished
    return eng, reqs


@pytest.mark.gpu
@pytest.mark.parametrize("cfg", [dict(hidden=128, heads=4, dtype=torch.float32),
                                 dict(hidden=256, heads=2, block_size=16, dtype=torch.bfloat16)], ids=["f32_d32", "bf16_d128"])
def test_decode_graphs_match_eager_engine(cfg):
    """The same engine run with decode-only steps replayed as CUDA graphs and run eagerly: same greedy tokens,
    same logits (padded batch sizes may pick another GEMM kernel: fp32 to 1e-4, bf16 to the bf16 tolerance)."""
    prompts = [list(range(3, 3 + n)) for n in (40, 9, 21, 33, 5)]
    kw = dict(cfg, token_budget=48, max_seqs=8)
    a, ra = _run("cuda", None, False, prompts, **kw)
    b, rb = _run("cuda", None, True, prompts, **kw)
    g = b._pa_b200["graphs"]
    assert a._pa_b200["graphs"] is None and g is not None
    assert g.replays >= 10 and g.captures >= 4, (g.replays, g.captures)      # decode buckets 1, 2, 4, 8 + mixed buckets
    assert any(k[2] for k in g.graphs), "no mixed chunked-prefill + decode step was replayed from a graph"
    tol = 1e-4 if cfg["dtype"] == torch.float32 else 6e-2
    same = 0
    for x, y in zip(ra, rb):
        same += x.output_token_ids == y.output_token_ids
        la, lb = a.logits_log[x.request_id], b.logits_log[y.request_id]
        if x.output_token_ids == y.output_token_ids:
            for u, v in zip(la, lb):
                assert torch.allclose(u, v, atol=tol, rtol=tol), (u - v).abs().max()
    assert same == len(ra) if cfg["dtype"] == torch.float32 else same >= len(ra) - 2
    # and against the CPU oracle engine (fp32: the golden side of the engine tests)
    if cfg["dtype"] == torch.float32:
        c, rc = _run("cpu", OracleOps, False, prompts, **kw)
        for x, y in zip(rc, rb):
            assert x.output_token_ids == y.output_token_ids
            for u, v in zip(c.logits_log[x.request_id], b.logits_log[y.request_id]):
                assert torch.allclose(u, v, atol=2e-4, rtol=1e-4), (u - v).abs().max()


@pytest.mark.gpu
def test_decode_graphs_survive_capacity_growth_and_uninstall():
    """More requests than the scheduler's nominal max_num_seqs are never scheduled, but a longer context than the
    initial table width is: the staging mirror is re-allocated, captured graphs are dropped and re-captured."""
    prompts = [list(range(1, 60)), list(range(5, 30))]
    eng = TinyEngine(device="cuda", hidden=128, heads=4, num_blocks=128, block_size=4, token_budget=64, max_seqs=2)
    engine_adapter.install(eng)
    eng._pa_b200["tables"]._alloc(2, 64, 8, 0)                    # narrow tables: 59 + 40 tokens need 25 blocks
    reqs = [eng.add_request(p, max_new_tokens=40) for p in prompts]
    eng.run()
    g = eng._pa_b200["graphs"]
    assert eng._pa_b200["tables"].W >= 25 and g.captures >= 2 and g.replays >= 30
    ref = TinyEngine(device="cpu", hidden=128, heads=4, num_blocks=128, block_size=4, token_budget=64, max_seqs=2)
    engine_adapter.install(ref, ops=OracleOps)
    rr = [ref.add_request(p, max_new_tokens=40) for p in prompts]
    ref.run()
    for x, y in zip(rr, reqs):
        assert x.output_token_ids == y.output_token_ids
    engine_adapter.uninstall(eng)
    assert "forward" not in eng.model.__dict__
