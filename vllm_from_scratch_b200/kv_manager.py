"""Prefix-cache wiring and preemption-by-swap for the reference engine loop (SURVEY.md section 8
row f4), layered on ``engine_adapter.install`` WITHOUT editing ``05_mini_vllm/mini_vllm.py``.

    engine = mini_vllm.MiniVLLM(..., device='cuda')
    engine_adapter.install(engine)
    mgr = kv_manager.install(engine, num_host_blocks=256)
    engine.generate(prompts)            # shared prefixes hit the cache; a full pool swaps instead of stalling

What the reference has and does not use (all lines relative to 05_mini_vllm/mini_vllm.py):
  * ``BlockAllocator.allocate(block_hash)`` / ``mark_cached`` (:62-104) implement a hash -> block
    prefix cache with ref counts, but ``Scheduler._ensure_blocks`` (:195-207) always calls
    ``allocate()`` without a hash and nothing ever calls ``mark_cached``: no request ever shares a
    block.  Here every FULL block of a request is registered under the chained hash of
    ``06_global_prefix_cache/global_kv_pool.py:471-500`` (``hash((parent, tokens, extra_key))``) once
    its KV has been written, and a new request adopts the longest cached chain of its prompt
    before its first schedule: those blocks enter its block table by reference and their tokens
    are never recomputed (``num_computed_tokens`` starts behind them).  Sound only with the
    adapter's attention, which reads history through the block table.
  * When a running request cannot get a block, the reference scheduler silently skips it
    (:231-233, :254-256) - with every request stuck, ``generate`` spins forever.  Swap is prose only
    (BLOG_PART3_V2.md:1600-1618).  Here the newest running requests are preempted: their KV blocks
    go to a pinned host pool (``swap_out``), the device blocks are freed, and the request comes
    back (``swap_in`` into whatever blocks are free then, block table rewritten) once there is
    room - no recomputation, bit-identical KV.

Pure host logic over the public attributes of the reference's ``Scheduler`` / ``BlockAllocator`` /
``Request`` (duck-typed, the reference is never imported).  The byte movement is ``ops.swap_out`` /
``ops.swap_in`` / ``ops.HostKVPool`` - by default this package's CUDA path (no CPU fallback).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional, Tuple


def compute_block_hash(parent_hash: int, token_ids: Tuple[int, ...], extra_key: Optional[int] = None) -> int:
    """``global_kv_pool.py:471-482``: a block's hash covers its tokens and, through the parent
    hash, the whole prefix before it."""
    return hash((parent_hash, token_ids, extra_key))


def compute_block_hashes(token_ids: List[int], block_size: int, extra_key: Optional[int] = None) -> List[int]:
    """Chained hashes of every FULL block of ``token_ids`` (``global_kv_pool.py:485-500``; seed 0)."""
    hashes, parent = [], 0
    for i in range(0, len(token_ids) - block_size + 1, block_size):
        parent = compute_block_hash(parent, tuple(token_ids[i:i + block_size]), extra_key)
        hashes.append(parent)
    return hashes


@dataclass
class KVStats:
    prefix_hit_blocks: int = 0
    prefix_hit_tokens: int = 0
    blocks_marked: int = 0
    preemptions: int = 0
    blocks_swapped_out: int = 0
    blocks_swapped_in: int = 0


@dataclass
class _Swapped:
    request: Any
    host_blocks: List[int]


@dataclass
class KVManager:
    engine: Any
    ops: Any
    host_k: Any
    host_v: Any
    enable_prefix_cache: bool = True
    enable_swap: bool = True
    stats: KVStats = field(default_factory=KVStats)
    swapped: List[_Swapped] = field(default_factory=list)
    _marked: Dict[int, Tuple[int, int]] = field(default_factory=dict)   # id(req) -> (blocks marked, parent hash)
    _looked_up: set = field(default_factory=set)
    _orig_schedule: Any = None
    _orig_update: Any = None

    # ---- prefix cache -----------------------------------------------------------------
    def _adopt_cached_prefix(self, req) -> None:
        """Before a request's first schedule: take the longest cached chain of its prompt's
        full blocks by reference (``allocate(block_hash)`` bumps the ref count and pulls the block
        out of the free queue, :69-75) and start computing behind it.  At least one prompt token
        is always left to compute so the step produces the request's logits."""
        sched = self.engine.scheduler
        alloc, bs = sched.allocator, sched.block_size
        self._looked_up.add(id(req))
        if req.block_table or req.num_computed_tokens:
            return
        hashes = compute_block_hashes(req.prompt_token_ids, bs)[:(req.prompt_len - 1) // bs]
        parent, n = 0, 0
        for h in hashes:
            if h not in alloc.hash_to_block:
                break
            req.block_table.append(alloc.allocate(block_hash=h))
            parent, n = h, n + 1
        if n:
            req.num_computed_tokens = n * bs
            self._marked[id(req)] = (n, parent)
            # the reference's first-schedule branch starts every waiting request at position 0
            # (:269-290); its running branch resumes at num_computed_tokens (:239-262) - so a request
            # that starts behind a cached prefix enters through `running`
            sched.waiting = [r for r in sched.waiting if r is not req]
            sched.running.append(req)
            self.stats.prefix_hit_blocks += n
            self.stats.prefix_hit_tokens += n * bs

    def _mark_full_blocks(self, req) -> None:
        """After a step: every block whose ``block_size`` tokens have all been fed (so their KV is in
        the cache) becomes a prefix-cache entry under its chained hash (``mark_cached``, :98-103)."""
        sched = self.engine.scheduler
        alloc, bs = sched.allocator, sched.block_size
        done, parent = self._marked.get(id(req), (0, 0))
        full = min(req.num_computed_tokens // bs, len(req.block_table))
        if full <= done:
            return
        tokens = list(req.prompt_token_ids) + list(req.output_token_ids)
        for i in range(done, full):
            parent = compute_block_hash(parent, tuple(tokens[i * bs:(i + 1) * bs]))
            alloc.mark_cached(req.block_table[i], parent)
            self.stats.blocks_marked += 1
        self._marked[id(req)] = (full, parent)

    # ---- preemption by swap -------------------------------------------------------------
    def _blocks_wanted(self, running) -> int:
        """Blocks the reference scheduler will ask for when it walks ``running`` this step
        (decode: one token; chunked prefill: what is left of the token budget, :225-262)."""
        sched = self.engine.scheduler
        bs, budget, want = sched.block_size, sched.max_num_batched_tokens, 0
        for r in running:
            if r.is_finished or budget < 1:
                continue
            n = min(r.prompt_len - r.num_computed_tokens, budget) if r.is_prefill else 1
            budget -= n
            want += max(0, math.ceil((r.num_computed_tokens + n) / bs) - len(r.block_table))
        return want

    def _caches(self):
        m = self.engine.model
        return m.k_cache, m.v_cache                       # [L, NB, BS, H, D] each (:544-552): block axis 1

    def _swap_out(self, req) -> None:
        sched = self.engine.scheduler
        n = min(math.ceil(req.num_computed_tokens / sched.block_size), len(req.block_table))
        host_ids = self.host_k.allocate(n)                # raises "OOM" when the swap space is full
        pairs = list(zip(req.block_table[:n], host_ids))
        kc, vc = self._caches()
        for dev, host in ((kc, self.host_k), (vc, self.host_v)):
            ev = self.ops.swap_out(dev, host.kv_cache, pairs, block_dim=1)
            if ev is not None:
                ev.wait()                                 # compute stream: do not overwrite before the copy
        for b in req.block_table:
            sched.allocator.free(b)
        req.block_table.clear()
        self._marked.pop(id(req), None)                   # its blocks keep their hashes; the table is rebuilt on return
        sched.running = [r for r in sched.running if r is not req]
        self.swapped.append(_Swapped(req, host_ids))
        self.stats.preemptions += 1
        self.stats.blocks_swapped_out += n

    def _swap_in(self, entry: _Swapped) -> bool:
        sched = self.engine.scheduler
        alloc = sched.allocator
        n = len(entry.host_blocks)
        if alloc.num_free < n:
            return False
        new_blocks = [alloc.allocate() for _ in range(n)]
        pairs = list(zip(entry.host_blocks, new_blocks))
        kc, vc = self._caches()
        for dev, host in ((kc, self.host_k), (vc, self.host_v)):
            ev = self.ops.swap_in(host.kv_cache, dev, pairs, block_dim=1)
            if ev is not None:
                ev.wait()                                 # attention must see the restored blocks
        self.host_k.free(entry.host_blocks)
        req = entry.request
        req.block_table[:] = new_blocks
        # re-register the restored full blocks so later requests can share them again
        self._marked[id(req)] = (0, 0)
        sched.running.insert(0, req)                      # oldest work first: it is not the next victim
        self.stats.blocks_swapped_in += n
        return True

    # ---- the wrapped scheduler entry points -------------------------------------------------
    def schedule(self):
        sched = self.engine.scheduler
        alloc = sched.allocator
        # swapped requests sit at the head of `waiting` between steps only so that
        # `has_unfinished` (:339-341) stays true; the reference scheduler must never see them
        parked = {id(e.request) for e in self.swapped}
        if parked:
            sched.waiting = [r for r in sched.waiting if id(r) not in parked]

        if self.enable_swap:
            # 1. bring swapped requests back, oldest first, while the running set keeps its blocks
            while self.swapped and len(sched.running) < sched.max_num_seqs:
                e = self.swapped[0]
                reserve = self._blocks_wanted(sched.running) + 1       # + the block its own next token may need
                if alloc.num_free < len(e.host_blocks) + reserve and sched.running:
                    break
                if not self._swap_in(e):
                    if not sched.running:
                        raise RuntimeError(f"OOM: need {len(e.host_blocks)} blocks to resume "
                                           f"{e.request.request_id}, have {alloc.num_free}")
                    break
                self.swapped.pop(0)
            # 2. not enough blocks for the running set's next step: preempt the newest requests
            while len(sched.running) > 1 and self._blocks_wanted(sched.running) > alloc.num_free:
                victim = next((r for r in reversed(sched.running) if not r.is_finished), None)
                if victim is None or victim is sched.running[0]:
                    break
                self._swap_out(victim)

        if self.enable_prefix_cache:
            for req in list(sched.waiting):
                if len(sched.running) >= sched.max_num_seqs:
                    break
                if id(req) not in self._looked_up:
                    self._adopt_cached_prefix(req)

        out = self._orig_schedule()
        if self.swapped:
            sched.waiting = [e.request for e in self.swapped] + sched.waiting
        return out

    def update(self, scheduled, next_token_ids):
        # finished requests free their blocks inside the reference update (:325-333); snapshot the
        # tables first so their full blocks can still be registered (a freed block keeps its hash
        # and is a cache candidate until it is reallocated, :92-96)
        tables = [(s.request, list(s.request.block_table)) for s in scheduled]
        ret = self._orig_update(scheduled, next_token_ids)
        if self.enable_prefix_cache:
            for req, table in tables:
                if req.is_finished:
                    live = req.block_table
                    req.block_table = table
                    self._mark_full_blocks(req)
                    req.block_table = live
                    self._marked.pop(id(req), None)
                    self._looked_up.discard(id(req))
                else:
                    self._mark_full_blocks(req)
        return ret


def _default_ops():
    from . import paged_attention
    return paged_attention


def install(engine: Any, num_host_blocks: int = 0, ops: Any = None, enable_prefix_cache: bool = True,
            enable_swap: bool = True) -> KVManager:
    """Wire the prefix cache and swap preemption into ``engine`` (a ``MiniVLLM`` with
    ``engine_adapter.install`` already applied - the reference attention cannot read shared or
    relocated blocks).  ``num_host_blocks`` sizes the pinned swap space (default: as many blocks as
    the device pool).  Idempotent; returns the manager (``.stats``, ``.swapped``)."""
    if getattr(engine, "_pa_b200", None) is None:
        raise RuntimeError("kv_manager.install: call engine_adapter.install(engine) first")
    mgr = getattr(engine, "_pa_b200_kv", None)
    if mgr is not None:
        return mgr
    ops = ops or engine._pa_b200["ops"]
    k_cache, v_cache = engine.model.k_cache, engine.model.v_cache
    nh = num_host_blocks or k_cache.shape[1]
    host_k = ops.HostKVPool(k_cache, nh, block_dim=1) if enable_swap else None
    host_v = ops.HostKVPool(v_cache, nh, block_dim=1) if enable_swap else None
    mgr = KVManager(engine, ops, host_k, host_v, enable_prefix_cache, enable_swap)
    sched = engine.scheduler
    mgr._orig_schedule, mgr._orig_update = sched.schedule, sched.update      # the adapter's metadata wrapper stays inside
    sched.schedule, sched.update = mgr.schedule, mgr.update
    engine._pa_b200_kv = mgr
    return mgr


def uninstall(engine: Any) -> None:
    mgr = getattr(engine, "_pa_b200_kv", None)
    if mgr is None:
        return
    if mgr.swapped:
        raise RuntimeError("kv_manager.uninstall: requests are still swapped out")
    engine.scheduler.schedule, engine.scheduler.update = mgr._orig_schedule, mgr._orig_update
    engine._pa_b200_kv = None
