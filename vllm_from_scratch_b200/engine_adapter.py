"""Engine call site: puts the B200 paged-attention path under the reference's
``05_mini_vllm/mini_vllm.py`` engine loop WITHOUT editing that file.

    import mini_vllm                                  # the reference module, unmodified
    from vllm_from_scratch_b200 import engine_adapter
    engine = mini_vllm.MiniVLLM(..., device='cuda')
    engine_adapter.install(engine)
    engine.generate(prompts)

What is replaced (reference lines, relative to 05_mini_vllm/mini_vllm.py):
  * step 3 of ``SelfAttention.forward`` - the flat KV scatter ``k_cache_flat[slot_mapping] = k``
    (:421-427) -> ``reshape_and_cache`` ('bshd' layout, the model's own buffers, no copy);
  * step 4 (:429-485) - per-request attention.  The reference's decode branch attends over flat
    slot 0 only and ignores the block table (acknowledged in its comments, :462-471), and its
    prefill branch ignores earlier chunks of a chunked prefill (:437-458).  Here every query token
    attends causally over its request's whole history, read THROUGH the block table:
    on tensor-core shapes (bf16, block_size 16, head_dim 64 / 128) the whole mixed step - prefill
    chunks and decode tokens - is ONE ``paged_attention_prefill_varlen`` launch; otherwise decode
    tokens -> one ``paged_attention_decode_batched`` launch for all decode requests of the step,
    prefill chunks -> ``paged_attention_prefill_chunk``.
What is kept: QKV projection, RoPE, output projection (the module's own weights and ``rope``),
``Scheduler``, ``BlockAllocator``, ``Sampler``, ``MiniTransformer`` - untouched.

The model's ``forward`` is not given block tables (:731-748), so ``install`` wraps
``engine.scheduler.schedule`` and keeps, per step, the "attention metadata" of what was
scheduled (block tables, chunk start positions, context lengths) - the side channel vLLM calls
``attn_metadata``.

Row f3 lives here too.  The reference rebuilds every tensor of a step from
Python lists (:729-739).  ``StepTables`` keeps the requests' block tables ON THE DEVICE across steps
(``int32 [2 * max_num_seqs, max_blocks]``, one row per live request), diffs each scheduled request's
``block_table`` against a host mirror and ships only the new block ids - together with the step's
tokens, positions, slots, context lengths and ``cu_q_lens`` - as ONE pinned staging copy per step;
``pa_block_table_step`` applies the delta and gathers the scheduled rows in scheduling order.  Every
device-side metadata tensor of ``StepMetadata`` is a view of that staging mirror at a fixed address, so
decode-only steps are replayed as CUDA graphs (one per batch-size bucket) captured over the reference
model's own modules.

The adapter is duck-typed on the reference's attribute names and never imports the reference.
``ops`` is the kernel namespace (default: this package's CUDA module, which has no CPU
fallback); tests inject an oracle-backed namespace to check the plumbing against the real
reference engine on CPU.
"""
from __future__ import annotations

import os
import types
from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional

import torch


@dataclass
class StepMetadata:
    """What the scheduler handed to the model this step, in scheduling order."""
    num_tokens: List[int] = field(default_factory=list)      # tokens fed per request
    start_pos: List[int] = field(default_factory=list)       # position of the first of them
    is_prefill: List[bool] = field(default_factory=list)
    block_tables: List[List[int]] = field(default_factory=list)   # the requests' LIVE lists (not copies)
    all_decode: bool = False                                 # every request of the step feeds one decode token
    # device-side forms: views of StepTables' staging mirror, shared by every layer
    decode_rows: Optional[torch.Tensor] = None               # int64 [Bd] token index of each decode request
    decode_block_tables: Optional[torch.Tensor] = None       # int32 [Bd, max_blocks]
    decode_ctx_lens: Optional[torch.Tensor] = None           # int32 [Bd] = position + 1
    prefill_chunks: List[tuple] = field(default_factory=list)   # (token offset, n, start_pos, int32 block-table row)
    # varlen form of the WHOLE step (prefill chunks and decode tokens alike) for the one-launch
    # tensor-core path: request i owns tokens [cu_q_lens[i], cu_q_lens[i + 1]) ending at ctx_lens[i]
    all_block_tables: Optional[torch.Tensor] = None          # int32 [B, max_blocks]
    cu_q_lens: Optional[torch.Tensor] = None                 # int32 [B + 1]
    ctx_lens: Optional[torch.Tensor] = None                  # int32 [B]
    max_q_len: int = 0
    # split route of a mixed step: the same varlen launch with the decode requests masked out
    # (context length 0 = skipped by the kernel) + one decode launch for them
    ctx_lens_prefill_only: Optional[torch.Tensor] = None     # int32 [B]
    decode_kv_tokens: int = 0                                # sum of the decode requests' context lengths
    # the step's model inputs as staged (graph replay reads these instead of the tensors step() built)
    token_ids: Optional[torch.Tensor] = None                 # int64 [T]
    positions: Optional[torch.Tensor] = None                 # int64 [T]
    slot_mapping: Optional[torch.Tensor] = None              # int64 [T]
    last_idx: Optional[torch.Tensor] = None                  # int64 [B] index of each request's last token
    # per-TOKEN form for shapes outside the tensor-core prefill kernel: token t is a decode "sequence" of length
    # position + 1 over its request's table, so a whole mixed step is ONE generic decode launch
    tok_block_tables: Optional[torch.Tensor] = None          # int32 [T, max_blocks]
    tok_ctx_lens: Optional[torch.Tensor] = None              # int32 [T]


class StepTables:
    """Persistent device block tables + ONE pinned staging copy per step (row f3).

    ``table`` int32 ``[2 * cap, W]`` holds one row per live request and is only ever touched by
    ``ops.block_table_step``.  ``build(scheduled)`` diffs each request's ``block_table`` list against
    the host mirror of its row, writes the (row, column, block id) delta and the step's flat arrays
    into the pinned staging buffer, issues one H2D copy and one ``block_table_step`` call, and returns
    a ``StepMetadata`` whose tensors are views of the device mirror.  Capacities grow on demand (the
    views then move, so captured graphs are dropped - ``generation`` tells)."""

    _I32 = ("hdr", "row_of", "gather_lens", "cu_q_lens", "ctx_pf")
    _I64 = ("decode_rows", "last_idx", "tokens", "positions", "slots")

    TOKEN_TABLE_LIMIT = 4 << 20     # entries of the per-token table (16 MiB); beyond it that form is switched off

    def __init__(self, device, ops, block_size: int, max_seqs: int = 16, max_tokens: int = 256, max_blocks: int = 64,
                 token_tables: bool = False):
        self.device = torch.device(device)
        self.ops = ops
        self.block_size = int(block_size)
        self.token_tables = bool(token_tables)
        self.generation = 0
        self.step_no = 0
        self.updates_sent = 0                                   # block-table entries shipped so far (tests, stats)
        self._alloc(max(1, int(max_seqs)), max(1, int(max_tokens)), max(1, int(max_blocks)), 0)

    # ---- storage ------------------------------------------------------------------------
    def _alloc(self, cap: int, tcap: int, width: int, ucap: int) -> None:
        self.cap, self.tcap, self.W = cap, tcap, width
        self.ucap = max(ucap, 1024, 8 * cap)
        self.rows = 2 * cap
        dev = self.device
        if self.token_tables and tcap * width > self.TOKEN_TABLE_LIMIT:
            self.token_tables = False
        ntok = tcap if self.token_tables else 0                 # gathered rows: [requests | decode requests | tokens]
        self.table = torch.zeros(self.rows, width, dtype=torch.int32, device=dev)
        self.step_tables = torch.zeros(2 * cap + ntok, width, dtype=torch.int32, device=dev)
        self.mirror: List[List[int]] = [[] for _ in range(self.rows)]
        self.key_of_row: List[Any] = [None] * self.rows
        self.row_of_key: Dict[Any, int] = {}
        self.last_used = [0] * self.rows
        self.free_rows = list(range(self.rows - 1, -1, -1))
        sizes = {"hdr": 4, "row_of": 2 * cap + ntok, "gather_lens": 2 * cap + ntok, "cu_q_lens": cap + 1, "ctx_pf": cap,
                 "decode_rows": cap, "last_idx": cap, "tokens": tcap, "positions": tcap, "slots": tcap}
        off, at = {}, 0
        for name in self._I32:
            off[name] = at
            at += 4 * sizes[name]
        at = (at + 7) // 8 * 8
        for name in self._I64:
            off[name] = at
            at += 8 * sizes[name]
        self._upd_off = at
        total = at + 12 * self.ucap
        cuda = dev.type == "cuda"
        self._host = [torch.zeros(total, dtype=torch.uint8, pin_memory=cuda) for _ in range(2 if cuda else 1)]
        self._host_done = [torch.cuda.Event() if cuda else None for _ in self._host]
        self._dev = torch.zeros(total, dtype=torch.uint8, device=dev)

        def views(buf, as_numpy):
            out = {}
            for name in self._I32 + self._I64:
                dt, es = (torch.int32, 4) if name in self._I32 else (torch.int64, 8)
                v = buf[off[name]:off[name] + es * sizes[name]].view(dt)
                out[name] = v.numpy() if as_numpy else v
            u = buf[self._upd_off:total].view(torch.int32).view(self.ucap, 3)
            out["updates"] = u.numpy() if as_numpy else u
            return out

        self._hv = [views(h, True) for h in self._host]
        self.dv = views(self._dev, False)
        for hv in self._hv:
            hv["row_of"][:] = -1
            hv["slots"][:] = -1
        self._prev = [(cap, cap, tcap)] * len(self._host)       # (B, Bd, T) last written into each host buffer
        self.generation += 1

    def _grow(self, batch: int, tokens: int, blocks: int, updates: int) -> None:
        cap, tcap, width, ucap = self.cap, self.tcap, self.W, self.ucap
        while cap < batch:
            cap *= 2
        while tcap < tokens:
            tcap *= 2
        while width < blocks:
            width *= 2
        while ucap < updates:
            ucap *= 2
        if self.device.type == "cuda":
            torch.cuda.current_stream(self.device).synchronize()    # nothing in flight still reads the old buffers
        self._alloc(cap, tcap, width, ucap)                     # mirrors are empty again: every row is re-sent

    def _take_row(self, key) -> int:
        if self.free_rows:
            row = self.free_rows.pop()
        else:                                                   # least recently scheduled row (never one of this step)
            row = min(range(self.rows), key=self.last_used.__getitem__)
            del self.row_of_key[self.key_of_row[row]]
        self.key_of_row[row] = key
        self.row_of_key[key] = row
        self.mirror[row] = []
        return row

    # ---- one step -------------------------------------------------------------------------
    def build(self, scheduled) -> StepMetadata:
        batch = len(scheduled)
        tokens = sum(len(s.tokens) for s in scheduled)
        blocks = max(len(s.request.block_table) for s in scheduled)
        if batch > self.cap or tokens > self.tcap or blocks > self.W:
            self._grow(batch, tokens, blocks, 0)
        md = self._fill(scheduled)
        if md is None:                                          # delta larger than the staging area: grow, re-send
            self._grow(batch, tokens, blocks, 2 * sum(len(s.request.block_table) for s in scheduled) + 1)
            md = self._fill(scheduled)
        return md

    def _fill(self, scheduled) -> Optional[StepMetadata]:
        self.step_no += 1
        step, cap, bs = self.step_no, self.cap, self.block_size
        which = step % len(self._host)
        if self._host_done[which] is not None:
            self._host_done[which].synchronize()                # the copy that last read this pinned buffer is done
        hv = self._hv[which]
        md = StepMetadata()
        rows, ctx, cu, last = [], [], [0], []
        dec_rows, dec_tok, dec_ctx, pf_ctx = [], [], [], []
        toks: List[int] = []
        pos: List[int] = []
        slots: List[int] = []
        upd: List[int] = []
        tok_rows: List[int] = []
        tok_lens: List[int] = []
        for sched in scheduled:
            req = sched.request
            n = len(sched.tokens)
            start = int(sched.positions[0])
            bt = req.block_table
            key = getattr(req, "request_id", None)
            if key is None:
                key = id(req)
            row = self.row_of_key.get(key)
            if row is None:
                row = self._take_row(key)
            elif self.last_used[row] == step:
                raise ValueError("engine_adapter: a request was scheduled twice in one step")
            self.last_used[row] = step
            mir = self.mirror[row]
            k = len(mir)
            if len(bt) < k or bt[:k] != mir:                    # not an append: blocks were replaced (swap-in, re-prefill)
                k = 0
                while k < len(mir) and k < len(bt) and bt[k] == mir[k]:
                    k += 1
                del mir[k:]
            for c in range(k, len(bt)):
                b = int(bt[c])
                upd += (row, c, b)
                mir.append(b)
            md.num_tokens.append(n)
            md.start_pos.append(start)
            md.is_prefill.append(bool(sched.is_prefill))
            md.block_tables.append(bt)
            rows.append(row)
            ctx.append(start + n)
            if sched.is_prefill:
                pf_ctx.append(start + n)
            else:
                if n != 1:
                    raise ValueError("a decode request feeds exactly one token per step")
                pf_ctx.append(0)
                dec_rows.append(row)
                dec_tok.append(cu[-1])
                dec_ctx.append(start + 1)
            cu.append(cu[-1] + n)
            last.append(cu[-1] - 1)
            toks += sched.tokens
            pos += sched.positions
            slots += sched.slot_mapping
            if self.token_tables:
                tok_rows += [row] * n
                tok_lens += range(start + 1, start + n + 1)
        nupd = len(upd) // 3
        if nupd > self.ucap:
            for row in rows:                                    # the delta was not shipped: forget it
                self.mirror[row] = []
            return None
        b, bd, t = len(rows), len(dec_rows), cu[-1]
        pb, pbd, pt = self._prev[which]
        # invariant of every host buffer: beyond the step's own entries row_of = -1, lengths 0, cu_q_lens = T,
        # tokens / positions 0, slots -1 (a padded request or token is a no-op for every kernel)
        hv["hdr"][0] = nupd
        hv["hdr"][1:4] = (b, bd, t)
        hv["row_of"][:b] = rows
        hv["row_of"][b:max(b, pb)] = -1
        hv["row_of"][cap:cap + bd] = dec_rows
        hv["row_of"][cap + bd:cap + max(bd, pbd)] = -1
        hv["gather_lens"][:b] = ctx
        hv["gather_lens"][b:max(b, pb)] = 0
        hv["gather_lens"][cap:cap + bd] = dec_ctx
        hv["gather_lens"][cap + bd:cap + max(bd, pbd)] = 0
        hv["cu_q_lens"][:b + 1] = cu
        hv["cu_q_lens"][b + 1:] = t
        hv["ctx_pf"][:b] = pf_ctx
        hv["ctx_pf"][b:max(b, pb)] = 0
        hv["decode_rows"][:bd] = dec_tok
        hv["last_idx"][:b] = last
        hv["last_idx"][b:max(b, pb)] = 0
        hv["tokens"][:t] = toks
        hv["tokens"][t:max(t, pt)] = 0
        hv["positions"][:t] = pos
        hv["positions"][t:max(t, pt)] = 0
        hv["slots"][:t] = slots
        hv["slots"][t:max(t, pt)] = -1
        if self.token_tables:
            hv["row_of"][2 * cap:2 * cap + t] = tok_rows
            hv["row_of"][2 * cap + t:2 * cap + max(t, pt)] = -1
            hv["gather_lens"][2 * cap:2 * cap + t] = tok_lens
            hv["gather_lens"][2 * cap + t:2 * cap + max(t, pt)] = 0
        if nupd:
            hv["updates"].reshape(-1)[:3 * nupd] = upd
        self._prev[which] = (b, bd, t)
        self.updates_sent += nupd
        nbytes = self._upd_off + 12 * nupd
        self._dev[:nbytes].copy_(self._host[which][:nbytes], non_blocking=True)
        if self._host_done[which] is not None:
            self._host_done[which].record()
        dv = self.dv
        self.ops.block_table_step(self.table, dv["updates"], dv["hdr"][0:1], dv["row_of"], dv["gather_lens"],
                                  self.step_tables, bs)
        md.all_decode = bd == b
        md.all_block_tables = self.step_tables[:b]
        md.cu_q_lens = dv["cu_q_lens"][:b + 1]
        md.ctx_lens = dv["gather_lens"][:b]
        md.max_q_len = max(md.num_tokens)
        md.token_ids, md.positions, md.slot_mapping = dv["tokens"][:t], dv["positions"][:t], dv["slots"][:t]
        md.last_idx = dv["last_idx"][:b]
        if self.token_tables:
            md.tok_block_tables = self.step_tables[2 * cap:2 * cap + t]
            md.tok_ctx_lens = dv["gather_lens"][2 * cap:2 * cap + t]
        if bd:
            md.decode_rows = dv["decode_rows"][:bd]
            md.decode_block_tables = self.step_tables[:b] if md.all_decode else self.step_tables[cap:cap + bd]
            md.decode_ctx_lens = dv["gather_lens"][:b] if md.all_decode else dv["gather_lens"][cap:cap + bd]
        at = 0
        for i, (n, start, pf) in enumerate(zip(md.num_tokens, md.start_pos, md.is_prefill)):
            if pf:
                md.prefill_chunks.append((at, n, start, self.step_tables[i]))
            at += n
        if bd and md.prefill_chunks:
            md.decode_kv_tokens = sum(dec_ctx)
            md.ctx_lens_prefill_only = dv["ctx_pf"][:b]
        return md

    def padded(self, bucket: int, tokens: int = 0, mixed: bool = False) -> StepMetadata:
        """The metadata of a padded step at the staging mirror's fixed addresses (what a CUDA graph is captured
        over): ``bucket`` requests, and ``tokens`` >= ``bucket`` tokens for a mixed step (0: a step whose requests
        all feed one token).  Requests / tokens past the real ones have context length 0, row -1, slot -1."""
        dv = self.dv
        cap = self.cap
        t = max(tokens, bucket) if mixed else bucket
        # any token split of the right total will do: only the totals shape the captured launches
        md = StepMetadata(num_tokens=[t - bucket + 1] + [1] * (bucket - 1), start_pos=[0] * bucket,
                          is_prefill=[mixed] + [False] * (bucket - 1), block_tables=[[]] * bucket, all_decode=not mixed)
        md.all_block_tables = self.step_tables[:bucket]
        md.ctx_lens = dv["gather_lens"][:bucket]
        md.cu_q_lens = dv["cu_q_lens"][:bucket + 1]
        md.max_q_len = t - bucket + 1 if mixed else 1
        if not mixed:
            md.decode_block_tables, md.decode_ctx_lens = md.all_block_tables, md.ctx_lens
            md.decode_rows = dv["decode_rows"][:bucket]
        elif self.token_tables:
            md.tok_block_tables = self.step_tables[2 * cap:2 * cap + t]
            md.tok_ctx_lens = dv["gather_lens"][2 * cap:2 * cap + t]
        else:
            md.max_q_len = t                                    # varlen launch: any chunk length up to the bucket
        md.token_ids, md.positions, md.slot_mapping = dv["tokens"][:t], dv["positions"][:t], dv["slots"][:t]
        md.last_idx = dv["last_idx"][:bucket]
        return md


# Mixed steps whose decode requests read at least this many KV bytes take the split route (decode
# kernel for the decode rows, varlen launch for the prefill chunks); below it the second launch costs
# more than it saves.
SPLIT_DECODE_BYTES = 32 << 20


def _rotate_half(x: torch.Tensor) -> torch.Tensor:
    half = x.shape[-1] // 2
    return torch.cat([-x[..., half:], x[..., :half]], dim=-1)


def _default_ops():
    from . import paged_attention
    return paged_attention


def _attention_forward(attn, engine, x, positions, kv_cache, slot_mapping, seq_lens, is_prefill):
    """Replacement for ``SelfAttention.forward`` (mini_vllm.py:394-487), bound per module."""
    state = engine._pa_b200
    md: StepMetadata = state["meta"]
    ops = state["ops"]
    if md is None or md.num_tokens != [int(s) for s in seq_lens] or md.is_prefill != [bool(p) for p in is_prefill]:
        raise RuntimeError("engine_adapter: the model was called with a batch the wrapped scheduler did not produce")
    t = x.shape[0]
    h, d = attn.num_heads, attn.head_dim

    k_cache, v_cache = kv_cache
    cache = (k_cache, v_cache)
    fused = getattr(ops, "rope_and_cache", None)
    if fused is not None and hasattr(attn.rope, "cos_cached"):
        # steps 1b-3 in one kernel: split, RoPE with the module's own tables, K / V slot write
        q = fused(attn.qkv(x), positions, attn.rope.cos_cached, attn.rope.sin_cached, cache, slot_mapping, h)
        return _attend(attn, ops, state, md, q, cache, t, h, d)

    # steps 1-2 of the reference forward, unchanged: QKV projection and RoPE (:408-419)
    q, k, v = attn.qkv(x).chunk(3, dim=-1)
    q, k, v = q.reshape(t, h, d), k.reshape(t, h, d), v.reshape(t, h, d)
    cos, sin = attn.rope(positions)
    cos, sin = cos.unsqueeze(1), sin.unsqueeze(1)
    q = (q * cos + _rotate_half(q) * sin).contiguous()
    k = (k * cos + _rotate_half(k) * sin).contiguous()
    v = v.contiguous()

    # step 3: KV slot write into the model's own [NB, BS, H, D] buffers (:421-427)
    ops.reshape_and_cache(k, v, cache, slot_mapping)
    return _attend(attn, ops, state, md, q, cache, t, h, d)


def _decode(ops, state, q, cache, block_tables, ctx_lens):
    ws = state.get("workspace")
    if ws is not None:           # adapter-owned split-K workspace: its address never changes under a captured graph
        return ops.paged_attention_decode_batched(q, cache, block_tables, ctx_lens, workspace=ws)
    return ops.paged_attention_decode_batched(q, cache, block_tables, ctx_lens)


def _attend(attn, ops, state, md: StepMetadata, q, cache, t, h, d):
    # step 4: attention through the block tables
    can_decode = hasattr(ops, "paged_attention_decode_batched")
    if md.all_decode and can_decode:
        # a decode-only step is the HBM-bound case: one decode launch for the whole batch, rows in place
        out = _decode(ops, state, q, cache, md.decode_block_tables, md.decode_ctx_lens)
        return attn.out_proj(out.reshape(t, h * d))
    varlen = getattr(ops, "paged_attention_prefill_varlen", None)
    if varlen is not None and ops.prefill_tensor_core_available(cache, h):
        kv_bytes = md.decode_kv_tokens * 2 * h * d * q.element_size()
        if md.ctx_lens_prefill_only is not None and can_decode and kv_bytes >= SPLIT_DECODE_BYTES:
            # a mixed step whose decode requests stream a lot of KV: a one-token query fills 1 / 32 of a
            # 128-row tensor-core tile and streams its KV well below the HBM rate there, so the decode rows
            # go to the HBM-bound decode kernel and the same varlen launch skips them (measured, eager
            # launches: 4 x 512-token chunks + 60 decode tokens at 2k context 250 us in one launch, 224 us
            # split - the split step is already bound by its four host-side launches)
            out = torch.empty((t, h, d), dtype=q.dtype, device=q.device)
            varlen(q, cache, md.all_block_tables, md.cu_q_lens, md.ctx_lens_prefill_only, md.max_q_len, out=out)
            o_dec = _decode(ops, state, q.index_select(0, md.decode_rows), cache, md.decode_block_tables,
                            md.decode_ctx_lens)
            out.index_copy_(0, md.decode_rows, o_dec)
            return attn.out_proj(out.reshape(t, h * d))
        # tensor-core shapes: the whole mixed step (prefill chunks + decode tokens) in ONE launch
        out = varlen(q, cache, md.all_block_tables, md.cu_q_lens, md.ctx_lens, md.max_q_len)
        return attn.out_proj(out.reshape(t, h * d))
    if md.tok_block_tables is not None and can_decode:
        # every other shape (fp32, toy head sizes): each token is a decode "sequence" of length position + 1 over its
        # request's table - the whole mixed step in ONE generic decode launch instead of one per prefill chunk
        out = _decode(ops, state, q, cache, md.tok_block_tables, md.tok_ctx_lens)
        return attn.out_proj(out.reshape(t, h * d))
    out = torch.empty((t, h, d), dtype=q.dtype, device=q.device)
    if md.decode_rows is not None:
        q_dec = q.index_select(0, md.decode_rows)
        o_dec = _decode(ops, state, q_dec, cache, md.decode_block_tables, md.decode_ctx_lens)
        out.index_copy_(0, md.decode_rows, o_dec)
    for offset, n, start, bt in md.prefill_chunks:
        out[offset:offset + n] = ops.paged_attention_prefill_chunk(q[offset:offset + n], cache, bt, start)
    return attn.out_proj(out.reshape(t, h * d))


# ---------------------------------------------------------------------------------------
# CUDA-graph replay of decode-only steps
# ---------------------------------------------------------------------------------------
_MODEL_ATTRS = ("embedding", "layers", "norm", "lm_head", "k_cache", "v_cache")


class DecodeGraphs:
    """Engine steps as CUDA graphs: decode-only steps, one graph per batch-size bucket (1, 2, 4, ...
    max_num_seqs), and mixed chunked-prefill + decode steps, one per (token bucket, request bucket).

    What is captured is the reference model's own forward (mini_vllm.py:592-613: embedding -> layers ->
    norm -> last-token rows -> lm_head) over the reference's own modules, with its inputs read from
    ``StepTables``' staging mirror (fixed addresses) and the last-token rows selected by a device index
    instead of a Python list (a list would be baked into the graph).  A replayed step therefore costs the
    host one staging fill, one H2D copy, the table step and ONE graph launch instead of ~15 launches per
    layer.  Padded requests and tokens (row -1, context 0, slot -1) are no-ops for every kernel.  A mixed
    step is capturable because it is ONE attention launch whatever its composition: the tcgen05 varlen kernel
    on tensor-core shapes, the generic decode kernel over per-token tables elsewhere."""

    def __init__(self, engine, state, mixed: bool = True):
        self.engine, self.state = engine, state
        self.graphs: Dict[tuple, tuple] = {}
        self.generation = -1
        self.replays = 0
        self.captures = 0
        self.mixed = mixed

    def usable(self, md: Optional[StepMetadata]) -> bool:
        tables: StepTables = self.state["tables"]
        if md is None or len(md.num_tokens) > tables.cap:
            return False
        if md.all_decode:
            return True
        if not self.mixed or sum(md.num_tokens) > tables.tcap:
            return False
        if tables.token_tables:
            return True
        # tensor-core shapes: one varlen launch - unless the decode rows stream enough KV for the split route
        h, d = self.state["heads"], self.state["head_dim"]
        return self.state["tc"] and md.decode_kv_tokens * 2 * h * d * 2 < SPLIT_DECODE_BYTES

    def _body(self, md: StepMetadata):
        m = self.engine.model
        x = m.embedding(md.token_ids)
        for i, layer in enumerate(m.layers):
            x = layer(x, md.positions, (m.k_cache[i], m.v_cache[i]), md.slot_mapping, md.num_tokens, md.is_prefill)
        return m.lm_head(m.norm(x).index_select(0, md.last_idx))

    @staticmethod
    def _bucket(n: int, cap: int) -> int:
        k = 1
        while k < n:
            k *= 2
        return min(k, cap)

    def run(self, md: StepMetadata) -> torch.Tensor:
        tables: StepTables = self.state["tables"]
        if self.generation != tables.generation:                # the staging mirror moved: captured addresses are stale
            self.graphs.clear()
            self.generation = tables.generation
        b = len(md.num_tokens)
        bk = self._bucket(b, tables.cap)
        if md.all_decode:
            key = (bk, bk, False)
            if key not in self.graphs:
                # first decode-only step (or the mirror moved): capture EVERY decode bucket now, so that no later
                # step stalls on a capture (~5 ms each, more when cuBLAS meets a new shape); padded requests are
                # no-ops and the step's own K / V rows are written with the same values by each warm-up pass
                k = 1
                while k < tables.cap:
                    if (k, k, False) not in self.graphs:
                        self._capture((k, k, False))
                    k *= 2
                if (tables.cap, tables.cap, False) not in self.graphs:
                    self._capture((tables.cap, tables.cap, False))
        else:
            key = (max(self._bucket(sum(md.num_tokens), tables.tcap), bk), bk, True)
            if key not in self.graphs:
                self._capture(key)
        graph, logits = self.graphs[key]
        graph.replay()
        self.replays += 1
        return logits[:b]

    def _capture(self, key: tuple):
        tables: StepTables = self.state["tables"]
        md_k = tables.padded(key[1], key[0], key[2])
        live = self.state["meta"]
        self.state["meta"] = md_k
        try:
            with torch.no_grad():
                side = torch.cuda.Stream()
                side.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(side):                   # warm-up outside capture (cuBLAS handles, tensor maps);
                    self._body(md_k)                            # idempotent: it writes this step's own K / V rows
                torch.cuda.current_stream().wait_stream(side)
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph):
                    logits = self._body(md_k)
        finally:
            self.state["meta"] = live
        self.graphs[key] = (graph, logits)
        self.captures += 1
        return graph, logits


def install(engine: Any, ops: Any = None, cuda_graphs: Optional[bool] = None) -> Any:
    """Route ``engine``'s attention through the paged-attention kernels.  Idempotent.

    ``cuda_graphs`` (default: on for CUDA engines whose model has the reference's attribute names, off
    with ``PA_B200_ENGINE_GRAPHS=0``): replay decode-only steps as CUDA graphs (``DecodeGraphs``)."""
    if getattr(engine, "_pa_b200", None) is not None:
        return engine
    ops = ops or _default_ops()
    layers = list(engine.model.layers)
    device = torch.device(engine.device)
    sched = engine.scheduler
    block_size = int(engine.model.k_cache.shape[2])
    a0 = layers[0].attn
    # shapes the tcgen05 varlen kernel does not take run a mixed step as one generic decode launch over per-token tables
    tc_fn = getattr(ops, "prefill_tensor_core_available", None)
    tc = False
    if tc_fn is not None and device.type == "cuda":
        tc = bool(tc_fn((engine.model.k_cache[0], engine.model.v_cache[0]), a0.num_heads))
    # capacities start at the scheduler's own limits (tokens capped: the staging grows on demand)
    tables = StepTables(device, ops, block_size, getattr(sched, "max_num_seqs", 16),
                        min(int(getattr(sched, "max_num_batched_tokens", 256)), 8192), 64,
                        token_tables=not tc and hasattr(ops, "paged_attention_decode_batched"))
    state = {"ops": ops, "meta": None, "orig_schedule": sched.schedule, "tables": tables, "workspace": None,
             "workspace_cap": 0, "orig_forwards": [layer.attn.forward for layer in layers], "graphs": None,
             "orig_model_forward": None, "tc": tc, "heads": a0.num_heads, "head_dim": a0.head_dim}
    engine._pa_b200 = state
    def new_workspace():
        state["workspace_cap"] = max(tables.cap, tables.tcap if tables.token_tables else 0, 64)
        state["workspace"] = ops.new_decode_workspace(state["workspace_cap"], a0.num_heads, a0.num_heads, a0.head_dim,
                                                      device)

    if device.type == "cuda" and hasattr(ops, "new_decode_workspace"):
        new_workspace()

    def schedule_with_metadata():
        scheduled = state["orig_schedule"]()
        state["meta"] = tables.build(scheduled) if scheduled else None
        if state["workspace"] is not None and max(tables.cap, tables.tcap if tables.token_tables else 0) > state["workspace_cap"]:
            new_workspace()
            tables.generation += 1                              # captured graphs hold the old workspace
        return scheduled

    sched.schedule = schedule_with_metadata
    for layer in layers:
        attn = layer.attn

        def forward(self, x, positions, kv_cache, slot_mapping, seq_lens, is_prefill, _engine=engine):
            return _attention_forward(self, _engine, x, positions, kv_cache, slot_mapping, seq_lens, is_prefill)

        attn.forward = types.MethodType(forward, attn)

    if cuda_graphs is None:
        cuda_graphs = os.environ.get("PA_B200_ENGINE_GRAPHS", "1") != "0"
    # graphs need the adapter-owned workspace (a shared one may be re-allocated under a captured graph)
    if cuda_graphs and state["workspace"] is not None and all(hasattr(engine.model, a) for a in _MODEL_ATTRS):
        graphs = DecodeGraphs(engine, state)
        state["graphs"] = graphs
        state["orig_model_forward"] = engine.model.forward

        def model_forward(token_ids, positions, slot_mapping, seq_lens, is_prefill):
            md = state["meta"]
            if graphs.usable(md) and md.num_tokens == [int(s) for s in seq_lens] \
                    and md.is_prefill == [bool(p) for p in is_prefill]:
                return graphs.run(md)
            return state["orig_model_forward"](token_ids, positions, slot_mapping, seq_lens, is_prefill)

        engine.model.forward = model_forward
    return engine


def uninstall(engine: Any) -> None:
    state = getattr(engine, "_pa_b200", None)
    if state is None:
        return
    engine.scheduler.schedule = state["orig_schedule"]
    for layer, fwd in zip(engine.model.layers, state["orig_forwards"]):
        layer.attn.forward = fwd
    if state["orig_model_forward"] is not None:
        del engine.model.forward                                # back to the class's forward
    engine._pa_b200 = None


def current_metadata(engine: Any) -> Optional[StepMetadata]:
    """The metadata of the step being executed (for tests and for swap / copy drivers)."""
    state = getattr(engine, "_pa_b200", None)
    return None if state is None else state["meta"]
