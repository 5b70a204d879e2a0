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

The adapter is duck-typed on the reference's attribute names and never imports the reference.
``ops`` is the kernel namespace (default: this package's CUDA module, which has no CPU
fallback); tests inject an oracle-backed namespace to check the plumbing against the real
reference engine on CPU.
"""
from __future__ import annotations

import types
from dataclasses import dataclass, field
from typing import Any, List, Optional

import torch


@dataclass
class StepMetadata:
    """What the scheduler handed to the model this step, in scheduling order."""
    num_tokens: List[int] = field(default_factory=list)      # tokens fed per request
    start_pos: List[int] = field(default_factory=list)       # position of the first of them
    is_prefill: List[bool] = field(default_factory=list)
    block_tables: List[List[int]] = field(default_factory=list)
    # device-side forms, built once per step and shared by every layer
    decode_rows: Optional[torch.Tensor] = None               # int64 [Bd] token index of each decode request
    decode_block_tables: Optional[torch.Tensor] = None       # int32 [Bd, max_blocks]
    decode_ctx_lens: Optional[torch.Tensor] = None           # int32 [Bd] = position + 1
    prefill_chunks: List[tuple] = field(default_factory=list)   # (token offset, n, start_pos, int32 block table)
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


def _build_metadata(scheduled, device) -> StepMetadata:
    md = StepMetadata()
    offset = 0
    dec_rows, dec_bts, dec_lens = [], [], []
    for sched in scheduled:
        n = len(sched.tokens)
        start = int(sched.positions[0])
        bt = [int(b) for b in sched.request.block_table]
        md.num_tokens.append(n)
        md.start_pos.append(start)
        md.is_prefill.append(bool(sched.is_prefill))
        md.block_tables.append(bt)
        if sched.is_prefill:
            md.prefill_chunks.append((offset, n, start, torch.tensor(bt, dtype=torch.int32).to(device)))
        else:
            if n != 1:
                raise ValueError("a decode request feeds exactly one token per step")
            dec_rows.append(offset)
            dec_bts.append(bt)
            dec_lens.append(start + 1)
        offset += n
    if dec_rows:
        width = max(len(b) for b in dec_bts)
        table = torch.zeros(len(dec_bts), max(width, 1), dtype=torch.int32)
        for i, b in enumerate(dec_bts):
            table[i, :len(b)] = torch.tensor(b, dtype=torch.int32)
        md.decode_rows = torch.tensor(dec_rows, dtype=torch.int64).to(device)
        md.decode_block_tables = table.to(device)
        md.decode_ctx_lens = torch.tensor(dec_lens, dtype=torch.int32).to(device)
    if md.num_tokens:
        width = max(max(len(b) for b in md.block_tables), 1)
        table = torch.zeros(len(md.block_tables), width, dtype=torch.int32)
        for i, b in enumerate(md.block_tables):
            table[i, :len(b)] = torch.tensor(b, dtype=torch.int32)
        cu = [0]
        for n in md.num_tokens:
            cu.append(cu[-1] + n)
        md.all_block_tables = table.to(device)
        md.cu_q_lens = torch.tensor(cu, dtype=torch.int32).to(device)
        md.ctx_lens = torch.tensor([s + n for s, n in zip(md.start_pos, md.num_tokens)], dtype=torch.int32).to(device)
        md.max_q_len = max(md.num_tokens)
        if dec_rows and md.prefill_chunks:
            md.decode_kv_tokens = sum(dec_lens)
            md.ctx_lens_prefill_only = torch.tensor([s + n if pf else 0 for s, n, pf in
                                                     zip(md.start_pos, md.num_tokens, md.is_prefill)],
                                                    dtype=torch.int32).to(device)
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
        return _attend(attn, ops, md, q, cache, t, h, d)

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
    return _attend(attn, ops, md, q, cache, t, h, d)


def _attend(attn, ops, md: StepMetadata, q, cache, t, h, d):
    # step 4: attention through the block tables
    varlen = getattr(ops, "paged_attention_prefill_varlen", None)
    if varlen is not None and ops.prefill_tensor_core_available(cache, h):
        decode = getattr(ops, "paged_attention_decode_batched", None)
        kv_bytes = md.decode_kv_tokens * 2 * h * d * q.element_size()
        if md.ctx_lens_prefill_only is not None and decode is not None and kv_bytes >= SPLIT_DECODE_BYTES:
            # a mixed step whose decode requests stream a lot of KV: a one-token query fills 1 / 32 of a
            # 128-row tensor-core tile and streams its KV well below the HBM rate there, so the decode rows
            # go to the HBM-bound decode kernel and the same varlen launch skips them (measured, eager
            # launches: 4 x 512-token chunks + 60 decode tokens at 2k context 250 us in one launch, 224 us
            # split - the split step is already bound by its four host-side launches)
            out = torch.empty((t, h, d), dtype=q.dtype, device=q.device)
            varlen(q, cache, md.all_block_tables, md.cu_q_lens, md.ctx_lens_prefill_only, md.max_q_len, out=out)
            o_dec = decode(q.index_select(0, md.decode_rows), cache, md.decode_block_tables, md.decode_ctx_lens)
            out.index_copy_(0, md.decode_rows, o_dec)
            return attn.out_proj(out.reshape(t, h * d))
        # tensor-core shapes: the whole mixed step (prefill chunks + decode tokens) in ONE launch
        out = varlen(q, cache, md.all_block_tables, md.cu_q_lens, md.ctx_lens, md.max_q_len)
        return attn.out_proj(out.reshape(t, h * d))
    out = torch.empty((t, h, d), dtype=q.dtype, device=q.device)
    if md.decode_rows is not None:
        q_dec = q.index_select(0, md.decode_rows)
        o_dec = ops.paged_attention_decode_batched(q_dec, cache, md.decode_block_tables, md.decode_ctx_lens)
        out.index_copy_(0, md.decode_rows, o_dec)
    for offset, n, start, bt in md.prefill_chunks:
        out[offset:offset + n] = ops.paged_attention_prefill_chunk(q[offset:offset + n], cache, bt, start)
    return attn.out_proj(out.reshape(t, h * d))


def install(engine: Any, ops: Any = None) -> Any:
    """Route ``engine``'s attention through the paged-attention kernels.  Idempotent."""
    if getattr(engine, "_pa_b200", None) is not None:
        return engine
    ops = ops or _default_ops()
    layers = list(engine.model.layers)
    state = {"ops": ops, "meta": None, "orig_schedule": engine.scheduler.schedule,
             "orig_forwards": [layer.attn.forward for layer in layers]}
    engine._pa_b200 = state
    device = engine.device

    def schedule_with_metadata():
        scheduled = state["orig_schedule"]()
        state["meta"] = _build_metadata(scheduled, device) if scheduled else None
        return scheduled

    engine.scheduler.schedule = schedule_with_metadata
    for layer in layers:
        attn = layer.attn

        def forward(self, x, positions, kv_cache, slot_mapping, seq_lens, is_prefill, _engine=engine):
            return _attention_forward(self, _engine, x, positions, kv_cache, slot_mapping, seq_lens, is_prefill)

        attn.forward = types.MethodType(forward, attn)
    return engine


def uninstall(engine: Any) -> None:
    state = getattr(engine, "_pa_b200", None)
    if state is None:
        return
    engine.scheduler.schedule = state["orig_schedule"]
    for layer, fwd in zip(engine.model.layers, state["orig_forwards"]):
        layer.attn.forward = fwd
    engine._pa_b200 = None


def current_metadata(engine: Any) -> Optional[StepMetadata]:
    """The metadata of the step being executed (for tests and for swap / copy drivers)."""
    state = getattr(engine, "_pa_b200", None)
    return None if state is None else state["meta"]
