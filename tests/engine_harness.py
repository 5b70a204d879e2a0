"""Test infrastructure for the engine call site (config C4).

Two things live here, neither of which is product code:

* ``OracleOps`` - the kernel namespace ``engine_adapter.install(engine, ops=...)`` expects,
  backed by the CPU oracle.  Used to check the adapter's plumbing against the REAL reference
  engine in the build container (CPU), and as the expected-value side of the GPU tests.
* ``TinyEngine`` - a small stand-in for ``05_mini_vllm/mini_vllm.py``'s ``MiniVLLM`` that exposes the
  same attribute names the adapter is duck-typed on (``scheduler.schedule`` returning objects with
  ``request.block_table / tokens / positions / is_prefill / slot_mapping``; ``model.layers[i].attn``
  with ``qkv / out_proj / rope / num_heads / head_dim``; ``model.k_cache / v_cache``
  ``[L, NB, BS, H, D]``; ``device``).  ``/root/reference`` does not exist on the GPU box, so the GPU
  tests drive the adapter with this engine: decode-first scheduling, chunked prefill under a token
  budget, optional shared prefix blocks, greedy sampling.  It is written from the reference's
  interface, not from its code.
"""
from __future__ import annotations

import os
import sys
from dataclasses import dataclass, field
from typing import Dict, List, Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import paged_attention_oracle as O  # noqa: E402


class OracleOps:
    """CPU oracle behind the adapter's ``ops`` interface ('bshd' caches only)."""

    @staticmethod
    def reshape_and_cache(key, value, kv_cache, slot_mapping):
        k_cache, v_cache = kv_cache
        O.reshape_and_cache_bshd(k_cache, v_cache, key, value, slot_mapping)

    block_table_step = staticmethod(O.block_table_step)

    @staticmethod
    def paged_attention_decode_batched(query, kv_cache, block_tables, seq_lens):
        k_cache, v_cache = kv_cache
        return O.decode_batched_bshd(query, k_cache, v_cache, block_tables, seq_lens,
                                     compute_dtype=torch.float64).to(query.dtype)

    @staticmethod
    def paged_attention_prefill_chunk(query, kv_cache, block_table, start_pos):
        k_cache, v_cache = kv_cache
        kv = torch.stack([k_cache.permute(0, 2, 1, 3), v_cache.permute(0, 2, 1, 3)])
        return O.chunked_prefill_hbsd(query, kv, block_table.tolist(), start_pos, k_cache.shape[1],
                                      compute_dtype=torch.float64).to(query.dtype)


    # swap space for kv_manager (CPU stand-ins of HostKVPool / swap_out / swap_in)
    class HostKVPool:
        def __init__(self, like, num_blocks, block_dim=None):
            bd = like.dim() - 4 if block_dim is None else block_dim
            shape = list(like.shape)
            shape[bd] = num_blocks
            self.kv_cache = torch.zeros(shape, dtype=like.dtype)
            self.free_blocks = list(range(num_blocks))

        def allocate(self, n):
            if len(self.free_blocks) < n:
                raise RuntimeError(f"OOM: need {n}, have {len(self.free_blocks)}")
            return [self.free_blocks.pop(0) for _ in range(n)]

        def free(self, ids):
            self.free_blocks.extend(ids)

    @staticmethod
    def swap_out(dev_cache, host_cache, pairs, block_dim=None):
        O.swap_blocks(dev_cache, host_cache, pairs, dev_cache.dim() - 4 if block_dim is None else block_dim)

    @staticmethod
    def swap_in(host_cache, dev_cache, pairs, block_dim=None):
        O.swap_blocks(host_cache, dev_cache, pairs, dev_cache.dim() - 4 if block_dim is None else block_dim)


# ---------------------------------------------------------------------------------------
# TinyEngine
# ---------------------------------------------------------------------------------------
@dataclass
class Req:
    request_id: str
    prompt_token_ids: List[int]
    max_new_tokens: int
    output_token_ids: List[int] = field(default_factory=list)
    num_computed_tokens: int = 0
    block_table: List[int] = field(default_factory=list)
    is_finished: bool = False

    @property
    def in_prefill(self) -> bool:
        return self.num_computed_tokens < len(self.prompt_token_ids)

    # the reference Request's names for the same things (kv_manager is duck-typed on them)
    is_prefill = in_prefill

    @property
    def prompt_len(self) -> int:
        return len(self.prompt_token_ids)


@dataclass
class Sched:
    request: Req
    tokens: List[int]
    positions: List[int]
    is_prefill: bool
    slot_mapping: List[int]


class Blocks:
    """Ref-counted free list; blocks may be shared between requests (prefix sharing)."""

    def __init__(self, n: int):
        self.free = list(range(n))
        self.refs = [0] * n

    def take(self) -> int:
        if not self.free:
            raise RuntimeError("OOM: no free KV blocks")
        b = self.free.pop(0)
        self.refs[b] = 1
        return b

    def share(self, b: int) -> int:
        self.refs[b] += 1
        return b

    def release(self, b: int) -> None:
        self.refs[b] -= 1
        if self.refs[b] == 0:
            self.free.append(b)


class HashedBlocks:
    """Allocator with the INTERFACE of the reference BlockAllocator (allocate(block_hash) ->
    id | None, free(id), mark_cached(id, hash), num_free, hash_to_block), written from that
    interface: free blocks are reused oldest-first and keep their hash until reused."""

    def __init__(self, n: int):
        self.ref_count = [0] * n
        self.free_order: List[int] = list(range(n))
        self.hash_to_block: Dict[int, int] = {}
        self.block_to_hash: Dict[int, int] = {}

    @property
    def num_free(self) -> int:
        return len(self.free_order)

    def allocate(self, block_hash: Optional[int] = None) -> Optional[int]:
        if block_hash is not None and block_hash in self.hash_to_block:
            b = self.hash_to_block[block_hash]
            self.ref_count[b] += 1
            if b in self.free_order:
                self.free_order.remove(b)
            return b
        if not self.free_order:
            return None
        b = self.free_order.pop(0)
        h = self.block_to_hash.pop(b, None)
        if h is not None:
            self.hash_to_block.pop(h, None)
        self.ref_count[b] = 1
        return b

    def free(self, b: int) -> None:
        if self.ref_count[b] <= 0:
            return
        self.ref_count[b] -= 1
        if self.ref_count[b] == 0:
            self.free_order.append(b)

    release = free                                     # TinyScheduler's name

    def mark_cached(self, b: int, block_hash: int) -> None:
        if block_hash not in self.hash_to_block:
            self.hash_to_block[block_hash] = b
            self.block_to_hash[b] = block_hash

    # TinyScheduler's names
    def take(self) -> Optional[int]:
        return self.allocate()


class TinyScheduler:
    def __init__(self, blocks: Blocks, block_size: int, max_seqs: int, token_budget: int):
        self.blocks, self.bs, self.max_seqs, self.budget = blocks, block_size, max_seqs, token_budget
        # the reference Scheduler's attribute names (kv_manager is duck-typed on them)
        self.allocator, self.block_size = blocks, block_size
        self.max_num_seqs, self.max_num_batched_tokens = max_seqs, token_budget
        self.waiting: List[Req] = []
        self.running: List[Req] = []
        self.paused: List[Req] = []          # swapped out: not scheduled until resumed

    def _grow(self, r: Req, upto_tokens: int) -> bool:
        while len(r.block_table) * self.bs < upto_tokens:
            b = self.blocks.take()
            if b is None:                            # reference behaviour: the request sits this step out
                return False
            r.block_table.append(b)
        return True

    def _slots(self, r: Req, start: int, n: int) -> List[int]:
        return [r.block_table[p // self.bs] * self.bs + p % self.bs for p in range(start, start + n)]

    def schedule(self) -> List[Sched]:
        out: List[Sched] = []
        budget = self.budget
        # decode first, then running prefills, then new requests
        order = [r for r in self.running if not r.in_prefill] + [r for r in self.running if r.in_prefill]
        fresh = []
        for r in order + self.waiting:
            if len(out) >= self.max_seqs or budget < 1:
                break
            if r in self.waiting:
                fresh.append(r)
            start = r.num_computed_tokens
            if r.in_prefill:
                n = min(len(r.prompt_token_ids) - start, budget)
                toks = r.prompt_token_ids[start:start + n]
            else:
                n, toks = 1, [r.output_token_ids[-1]]
            if not self._grow(r, start + n):
                if r in fresh:
                    fresh.remove(r)
                continue
            out.append(Sched(r, toks, list(range(start, start + n)), r.in_prefill, self._slots(r, start, n)))
            budget -= n
        for r in fresh:
            self.waiting.remove(r)
            self.running.append(r)
        return out

    def update(self, scheduled: List[Sched], next_tokens: List[int]) -> None:
        for s, tok in zip(scheduled, next_tokens):
            r = s.request
            r.num_computed_tokens += len(s.tokens)
            if r.in_prefill:
                continue
            r.output_token_ids.append(int(tok))
            if len(r.output_token_ids) >= r.max_new_tokens:
                r.is_finished = True
                for b in r.block_table:
                    self.blocks.release(b)
                self.running.remove(r)

    @property
    def has_unfinished(self) -> bool:
        return bool(self.waiting or self.running or self.paused)


class Rope(nn.Module):
    def __init__(self, dim: int, max_len: int = 4096, base: float = 10000.0):
        super().__init__()
        inv = 1.0 / (base ** (torch.arange(0, dim, 2).float() / dim))
        ang = torch.outer(torch.arange(max_len).float(), inv)
        ang = torch.cat([ang, ang], dim=-1)
        self.register_buffer("cos_cached", ang.cos())
        self.register_buffer("sin_cached", ang.sin())

    def forward(self, positions):
        return self.cos_cached[positions], self.sin_cached[positions]


class Attn(nn.Module):
    """Holds the weights the adapter uses.  Its own forward is deliberately absent: the engine
    only works with the adapter installed (the reference's own forward is what gets replaced)."""

    def __init__(self, hidden: int, heads: int):
        super().__init__()
        self.num_heads, self.head_dim = heads, hidden // heads
        self.qkv = nn.Linear(hidden, 3 * hidden, bias=False)
        self.out_proj = nn.Linear(hidden, hidden, bias=False)
        self.rope = Rope(self.head_dim)

    def forward(self, *a, **k):
        raise RuntimeError("TinyEngine needs engine_adapter.install(engine)")


class Layer(nn.Module):
    def __init__(self, hidden: int, heads: int, inter: int):
        super().__init__()
        self.attn = Attn(hidden, heads)
        self.up, self.down = nn.Linear(hidden, inter, bias=False), nn.Linear(inter, hidden, bias=False)
        self.n1, self.n2 = nn.LayerNorm(hidden), nn.LayerNorm(hidden)

    def forward(self, x, positions, kv_cache, slot_mapping, seq_lens, is_prefill):
        x = x + self.attn(self.n1(x), positions, kv_cache, slot_mapping, seq_lens, is_prefill)
        return x + self.down(F.gelu(self.up(self.n2(x))))


class TinyModel(nn.Module):
    def __init__(self, vocab, hidden, layers, heads, inter, num_blocks, block_size):
        super().__init__()
        self.emb = nn.Embedding(vocab, hidden)
        self.layers = nn.ModuleList([Layer(hidden, heads, inter) for _ in range(layers)])
        self.norm = nn.LayerNorm(hidden)
        self.head = nn.Linear(hidden, vocab, bias=False)
        hd = hidden // heads
        self.register_buffer("k_cache", torch.zeros(layers, num_blocks, block_size, heads, hd))
        self.register_buffer("v_cache", torch.zeros(layers, num_blocks, block_size, heads, hd))

    # the reference MiniTransformer's names for the same modules (engine_adapter.DecodeGraphs is duck-typed on them)
    embedding = property(lambda self: self.emb)
    lm_head = property(lambda self: self.head)

    def forward(self, token_ids, positions, slot_mapping, seq_lens, is_prefill):
        x = self.emb(token_ids)
        for i, layer in enumerate(self.layers):
            x = layer(x, positions, (self.k_cache[i], self.v_cache[i]), slot_mapping, seq_lens, is_prefill)
        last = torch.tensor(seq_lens).cumsum(0) - 1
        return self.head(self.norm(x)[last.to(x.device)])


class TinyEngine:
    def __init__(self, vocab=200, hidden=128, layers=2, heads=4, inter=256, num_blocks=64, block_size=16,
                 max_seqs=8, token_budget=48, device="cpu", dtype=torch.float32, seed=0, hashed_blocks=False):
        torch.manual_seed(seed)
        self.device = torch.device(device)
        self.block_size = block_size
        self.blocks = HashedBlocks(num_blocks) if hashed_blocks else Blocks(num_blocks)
        self.scheduler = TinyScheduler(self.blocks, block_size, max_seqs, token_budget)
        self.model = TinyModel(vocab, hidden, layers, heads, inter, num_blocks, block_size).to(self.device).to(dtype)
        self.logits_log: Dict[str, List[torch.Tensor]] = {}
        self._n = 0

    def add_request(self, prompt: List[int], max_new_tokens: int, shared_prefix_of: Optional[Req] = None,
                    shared_tokens: int = 0) -> Req:
        r = Req(f"req-{self._n}", list(prompt), max_new_tokens)
        self._n += 1
        if shared_prefix_of is not None:
            # prefix-cache hit: reuse the donor's first blocks, skip their tokens
            nshare = shared_tokens // self.block_size
            r.block_table = [self.blocks.share(b) for b in shared_prefix_of.block_table[:nshare]]
            r.num_computed_tokens = nshare * self.block_size
        self.scheduler.waiting.append(r)
        return r

    @torch.no_grad()
    def step(self):
        scheduled = self.scheduler.schedule()
        if not scheduled:
            return []
        toks = [t for s in scheduled for t in s.tokens]
        pos = [p for s in scheduled for p in s.positions]
        slots = [x for s in scheduled for x in s.slot_mapping]
        dev = self.device
        logits = self.model(torch.tensor(toks, device=dev), torch.tensor(pos, device=dev),
                            torch.tensor(slots, dtype=torch.long, device=dev),
                            [len(s.tokens) for s in scheduled], [s.is_prefill for s in scheduled])
        nxt = logits.float().argmax(-1).tolist()
        for s, row in zip(scheduled, logits):
            if not s.is_prefill or s.positions[-1] + 1 == len(s.request.prompt_token_ids):
                self.logits_log.setdefault(s.request.request_id, []).append(row.float().cpu())
        self.scheduler.update(scheduled, nxt)
        return scheduled

    def run(self, max_steps=10_000):
        n = 0
        while self.scheduler.has_unfinished and n < max_steps:
            if not self.step() and not self.scheduler.paused:
                break
            n += 1
        return n
