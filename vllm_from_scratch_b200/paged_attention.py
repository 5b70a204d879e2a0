"""B200 drop-in for the reference's ``01_paged_attention/paged_attention.py``.

Same module-level names and signatures as the reference (``BLOCK_SIZE``, ``KVCachePool``,
``compute_slot_mapping``, ``paged_attention_decode``, ``paged_attention_prefill`` -
paged_attention.py:12-136), plus the batched / cache-op entry points the metric needs
(``paged_attention_decode_batched``, ``reshape_and_cache``, ``copy_blocks``, ``swap_out``,
``swap_in``, ``HostKVPool``).  All tensor work is done by hand-written sm_100a kernels in
``libpa_b200.so`` (C ABI: include/pa_b200.h) on the caller's current CUDA stream.  Torch is
used for device memory and streams only.  There is NO CPU fallback: CPU tensors raise.

Cache layouts (described to the kernels by strides, never transposed):
  'hbsd'  one tensor ``kv_cache[2, NB, Hkv, BS, D]``         (paged_attention.py:26-30, :69)
  'bshd'  a pair ``(k_cache, v_cache)`` each ``[NB, BS, H, D]`` (05_mini_vllm/mini_vllm.py:544-552)
"""
from __future__ import annotations

import ctypes
import itertools
from typing import List, Optional, Sequence, Tuple, Union

import numpy as np
import torch

from . import _lib
from ._lib import PaKvCache

BLOCK_SIZE = 16  # tokens per block (paged_attention.py:12)

KvCacheArg = Union[torch.Tensor, Tuple[torch.Tensor, torch.Tensor]]

_DTYPES = {torch.bfloat16: _lib.PA_BF16, torch.float32: _lib.PA_F32}


# ---------------------------------------------------------------------------------------
# helpers
# ---------------------------------------------------------------------------------------
def _require_cuda(t: torch.Tensor, name: str) -> None:
    if not isinstance(t, torch.Tensor):
        raise TypeError(f"{name} must be a torch.Tensor, got {type(t).__name__}")
    if not t.is_cuda:
        raise RuntimeError(
            f"{name} is on {t.device}: the B200 paged-attention path has no CPU fallback; "
            "allocate the tensor on a CUDA device")


def _require_cuda_or_pinned(t: torch.Tensor, name: str) -> None:
    """Device tensors, or PINNED host tensors: under unified addressing the kernels read / write
    page-locked host memory in place over PCIe (zero-copy), which saves a staging copy and its
    launch latency for small per-step operands (the decode query and output)."""
    if not isinstance(t, torch.Tensor):
        raise TypeError(f"{name} must be a torch.Tensor, got {type(t).__name__}")
    if not t.is_cuda and not t.is_pinned():
        raise RuntimeError(
            f"{name} is pageable host memory: the B200 paged-attention path has no CPU fallback; "
            "allocate it on a CUDA device (or pin it for zero-copy access)")


def _stream_ptr(stream: Optional[torch.cuda.Stream] = None) -> int:
    return (stream or torch.cuda.current_stream()).cuda_stream


class _DeviceOf:
    """Make the tensor's device current for the duration of a launch (cheap when it already is)."""

    def __init__(self, t: torch.Tensor):
        self.idx = t.device.index
        self.prev = None

    def __enter__(self):
        cur = torch.cuda.current_device()
        if self.idx is not None and cur != self.idx:
            self.prev = cur
            torch.cuda.set_device(self.idx)

    def __exit__(self, *exc):
        if self.prev is not None:
            torch.cuda.set_device(self.prev)


def _describe_cache(kv_cache: KvCacheArg) -> Tuple[PaKvCache, torch.Tensor]:
    """Build the strided PaKvCache descriptor for either layout; returns (desc, k_plane)."""
    if isinstance(kv_cache, (tuple, list)):
        k, v = kv_cache
        _require_cuda(k, "k_cache")
        _require_cuda(v, "v_cache")
        if k.dim() != 4 or v.shape != k.shape or v.stride() != k.stride() or v.dtype != k.dtype:
            raise ValueError("'bshd' cache must be a (k_cache, v_cache) pair of identical [NB, BS, H, D] tensors")
        nb, bs, h, d = k.shape
        sb, ss, sh, sd = k.stride()
    else:
        _require_cuda(kv_cache, "kv_cache")
        if kv_cache.dim() != 5 or kv_cache.shape[0] != 2:
            raise ValueError("'hbsd' cache must be kv_cache[2, NB, Hkv, BS, D]")
        k, v = kv_cache[0], kv_cache[1]
        nb, h, bs, d = k.shape
        sb, sh, ss, sd = k.stride()
    if sd != 1 and d > 1:
        raise ValueError("the head_dim axis of the KV cache must be contiguous")
    if k.dtype not in _DTYPES:
        raise TypeError(f"unsupported KV cache dtype {k.dtype} (bf16 / fp32 only)")
    desc = PaKvCache(k.data_ptr(), v.data_ptr(), sb, sh, ss, nb, h, bs, d, _DTYPES[k.dtype], 0)
    return desc, k


def _split_cache(kv_cache: KvCacheArg):
    """(k_cache, v_cache) pair -> the two tensors; a single [2, NB, Hkv, BS, D] tensor -> (tensor, None)."""
    if isinstance(kv_cache, (tuple, list)):
        return kv_cache[0], kv_cache[1]
    return kv_cache, None


_workspaces = {}


def _workspace(device: torch.device, nbytes: int) -> torch.Tensor:
    """Per-device scratch for split-K partials, grown on demand (stream-ordered reuse: callers
    that launch decode concurrently on several streams must pass their own ``workspace``).
    Allocated zeroed: the kernels keep the counter header at its start zero between launches."""
    ws = _workspaces.get(device.index)
    if ws is None or ws.numel() < nbytes:
        ws = torch.zeros(max(nbytes, 2 << 20), dtype=torch.uint8, device=device)
        _workspaces[device.index] = ws
    return ws


def new_decode_workspace(batch: int, num_q_heads: int, num_kv_heads: int, head_dim: int,
                         device) -> torch.Tensor:
    """A private, correctly initialised workspace for ``paged_attention_decode_batched(workspace=...)``
    (zeroed counter header, include/pa_b200.h)."""
    return torch.zeros(decode_workspace_bytes(batch, num_q_heads, num_kv_heads, head_dim), dtype=torch.uint8,
                       device=device)


def _as_block_table(block_table, device) -> torch.Tensor:
    if isinstance(block_table, torch.Tensor):
        bt = block_table.to(device=device, dtype=torch.int32)
    else:
        bt = torch.tensor(list(block_table), dtype=torch.int32).to(device, non_blocking=False)
    return bt.contiguous()


# ---------------------------------------------------------------------------------------
# block pool (paged_attention.py:15-51)
# ---------------------------------------------------------------------------------------
class BlockFifo:
    """Host-side free list with the reference's observable order (paged_attention.py:31-46):
    FIFO pop from the head, freed blocks re-appended at the tail when their count hits zero."""

    def __init__(self, num_blocks: int):
        self.ref_counts = [0] * num_blocks
        self.free_blocks: List[int] = list(range(num_blocks))

    def allocate(self, num_blocks: int) -> List[int]:
        if len(self.free_blocks) < num_blocks:
            raise RuntimeError(f"OOM: need {num_blocks}, have {len(self.free_blocks)}")
        allocated = self.free_blocks[:num_blocks]
        del self.free_blocks[:num_blocks]
        for b in allocated:
            self.ref_counts[b] = 1
        return allocated

    def free(self, block_ids: Sequence[int]) -> None:
        for b in block_ids:
            self.ref_counts[b] -= 1
            if self.ref_counts[b] == 0:
                self.free_blocks.append(b)


class KVCachePool(BlockFifo):
    """Physical block pool ``kv_cache[2, L, NB, Hkv, BS, D]`` in HBM (paged_attention.py:15-51).

    Same constructor signature as the reference; ``device`` must be a CUDA device (the
    reference's default ``'cpu'`` raises here - no CPU fallback)."""

    def __init__(self, num_blocks: int, block_size: int, num_kv_heads: int, head_dim: int,
                 num_layers: int, dtype=torch.float32, device='cpu'):
        dev = torch.device(device)
        if dev.type != 'cuda':
            raise RuntimeError(
                f"KVCachePool(device={device!r}): the B200 paged-attention path has no CPU fallback; "
                "pass device='cuda'")
        if dtype not in _DTYPES:
            raise TypeError(f"unsupported KV cache dtype {dtype} (bf16 / fp32 only)")
        super().__init__(num_blocks)
        self.num_blocks = num_blocks
        self.block_size = block_size
        self.num_kv_heads = num_kv_heads
        self.head_dim = head_dim
        self.num_layers = num_layers
        self.kv_cache = torch.zeros(2, num_layers, num_blocks, num_kv_heads, block_size, head_dim,
                                    dtype=dtype, device=dev)

    def layer(self, layer_idx: int) -> torch.Tensor:
        """The per-layer view ``[2, NB, Hkv, BS, D]`` the attention functions take."""
        return self.kv_cache[:, layer_idx]

    def write_kv(self, layer_idx: int, block_id: int, slot_in_block: int,
                 key: torch.Tensor, value: torch.Tensor):
        """Single-token KV write (paged_attention.py:48-51); key/value ``[Hkv, D]``."""
        slot = torch.tensor([block_id * self.block_size + slot_in_block], dtype=torch.int64).to(self.kv_cache.device)
        reshape_and_cache(key.unsqueeze(0), value.unsqueeze(0), self.layer(layer_idx), slot)


# ---------------------------------------------------------------------------------------
# slot mapping (paged_attention.py:54-64)
# ---------------------------------------------------------------------------------------
def compute_slot_mapping(block_table: List[int], seq_len: int, block_size: int) -> List[int]:
    """``slot[i] = block_table[i // BS] * BS + i % BS`` - host integer logic, as in the
    reference (the scheduler's lists stay in Python; the device form is below)."""
    return [block_table[i // block_size] * block_size + (i % block_size) for i in range(seq_len)]


def compute_slot_mapping_device(block_table: torch.Tensor, start_pos: int, num_tokens: int,
                                block_size: int) -> torch.Tensor:
    """Device form: int32 block table on the GPU -> int64 slots for positions
    ``[start_pos, start_pos + num_tokens)`` (mini_vllm.py:209-219)."""
    _require_cuda(block_table, "block_table")
    bt = block_table.to(torch.int32).contiguous()
    need = (start_pos + num_tokens + block_size - 1) // block_size
    if bt.numel() < need:
        raise IndexError(f"block_table has {bt.numel()} entries, positions up to {start_pos + num_tokens} need {need}")
    slots = torch.empty(num_tokens, dtype=torch.int64, device=bt.device)
    with _DeviceOf(bt):
        rc = _lib.load().pa_slot_mapping(bt.data_ptr(), start_pos, num_tokens, block_size, slots.data_ptr(),
                                         _stream_ptr())
    _lib.check(rc, "pa_slot_mapping")
    return slots


def block_table_step(table: torch.Tensor, updates: Optional[torch.Tensor], num_updates: Optional[torch.Tensor],
                     row_of: Optional[torch.Tensor], ctx_lens: Optional[torch.Tensor],
                     step_table: Optional[torch.Tensor], block_size: int) -> None:
    """One engine step of the persistent device block tables (``pa_block_table_step``; replaces the
    per-step rebuild mini_vllm.py:729-739).  ``table`` int32 ``[rows, W]`` stays on the device across
    steps; ``updates`` int32 ``[cap, 3]`` holds the scheduler's delta as (row, column, block id) triples
    of which the first ``num_updates[0]`` (a DEVICE int32) are applied; then
    ``step_table[j, :n] = table[row_of[j], :n]`` with ``n = ceil(ctx_lens[j] / block_size)`` for every
    ``j < len(row_of)`` (``row_of[j] < 0`` leaves row j alone).  No host sync; graph-capturable."""
    _lib.load_ext().block_table_step(table, updates, num_updates, row_of, ctx_lens, step_table, int(block_size))


# ---------------------------------------------------------------------------------------
# KV slot write
# ---------------------------------------------------------------------------------------
def reshape_and_cache(key: torch.Tensor, value: torch.Tensor, kv_cache: KvCacheArg,
                      slot_mapping: torch.Tensor) -> None:
    """Scatter new K/V rows ``[T, Hkv, D]`` into the paged cache by slot (slot = block * BS +
    offset; negative = skip).  Bit-exact copy.  Replaces paged_attention.py:48-51, :119-123
    and mini_vllm.py:424-427."""
    ca, cb = _split_cache(kv_cache)
    _lib.load_ext().reshape_and_cache(key, value, ca, cb, slot_mapping)


def rope_and_cache(qkv: torch.Tensor, positions: torch.Tensor, cos_cached: torch.Tensor, sin_cached: torch.Tensor,
                   kv_cache: KvCacheArg, slot_mapping: torch.Tensor, num_q_heads: int,
                   q_out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Fused RoPE + QKV split + KV slot write (mini_vllm.py:410-427 in one kernel).

    qkv ``[T, (Hq + 2 Hkv) * D]`` straight from the QKV projection; ``cos_cached`` / ``sin_cached``
    ``[max_len, D]`` are the reference ``RotaryEmbedding``'s own buffers.  Returns rotated Q
    ``[T, Hq, D]``; rotated K and plain V land in their cache slots.  Bit-exact against the
    reference's op-by-op torch evaluation in the same dtype."""
    ca, cb = _split_cache(kv_cache)
    return _lib.load_ext().rope_and_cache(qkv, positions, cos_cached, sin_cached, ca, cb, slot_mapping, int(num_q_heads), q_out)


# ---------------------------------------------------------------------------------------
# decode
# ---------------------------------------------------------------------------------------
def decode_workspace_bytes(batch: int, num_q_heads: int, num_kv_heads: int, head_dim: int) -> int:
    return int(_lib.load().pa_decode_workspace_bytes(batch, num_q_heads, num_kv_heads, head_dim))


def paged_attention_decode_batched(query: torch.Tensor, kv_cache: KvCacheArg, block_tables: torch.Tensor,
                                   seq_lens: torch.Tensor, out: Optional[torch.Tensor] = None,
                                   workspace: Optional[torch.Tensor] = None, peers=None) -> torch.Tensor:
    """Batched decode: for every sequence b, ``softmax(q_b K_b^T / sqrt(D)) V_b`` over the
    ``seq_lens[b]`` tokens reached through ``block_tables[b]`` - i.e. the reference function
    (paged_attention.py:67-101) looped over the batch: ONE launch on the bf16 fast path (the split-K
    kernel merges its own partials), two (split-K + reduce) on the generic fp32 / odd-shape path.

    query ``[B, Hq, D]``; block_tables int32 ``[B, max_blocks]`` and seq_lens int32 ``[B]`` on the
    device (no host sync on the launch path); entries past ``ceil(seq_len / BS)`` are never read;
    ``seq_len == 0`` gives zeros.  Output dtype = query's, on the cache's device unless ``out`` is
    given.  ``query`` / ``out`` may also be PINNED host tensors (read / written in place over PCIe).
    A caller-provided ``workspace`` must come from ``new_decode_workspace`` (zero-initialised
    counter header).  ``peers`` (a ``_lib.PaPeerGather``, built by ``sharding.PeerGather``) fuses the
    KV-head-sharded output gather into the kernel: rows are also stored into the remote ranks'
    gather buffers over NVLink and arrival counters are bumped (``pa_decode_gather``)."""
    ca, cb = _split_cache(kv_cache)
    ext = _lib.load_ext()
    if workspace is None:
        # both layouts keep the KV heads on axis 2 ([2, NB, Hkv, BS, D] and [NB, BS, H, D]); malformed arguments
        # are reported by the extension's own checks
        need = ext.decode_workspace_bytes(query.shape[0], query.shape[1], ca.shape[2], query.shape[-1]) \
            if query.dim() == 3 and ca.dim() >= 3 else 0
        workspace = _workspace(ca.device, need)
    return ext.decode(query, ca, cb, block_tables, seq_lens, out, workspace, 0 if peers is None else ctypes.addressof(peers))


def paged_attention_decode(query: torch.Tensor, kv_cache: torch.Tensor, block_table: List[int], seq_len: int,
                           head_dim: int, num_q_heads: int, num_kv_heads: int,
                           block_size: int = BLOCK_SIZE) -> torch.Tensor:
    """Single-sequence decode with the reference's exact signature (paged_attention.py:67-76):
    query ``[Hq, D]``, kv_cache ``[2, NB, Hkv, BS, D]``, block_table a Python list -> ``[Hq, D]``."""
    _require_cuda(query, "query")
    _require_cuda(kv_cache, "kv_cache")
    if query.shape != (num_q_heads, head_dim):
        raise ValueError(f"query must be [{num_q_heads}, {head_dim}]")
    if kv_cache.dim() != 5 or kv_cache.shape[2] != num_kv_heads or kv_cache.shape[3] != block_size \
            or kv_cache.shape[4] != head_dim:
        raise ValueError("kv_cache must be [2, NB, num_kv_heads, block_size, head_dim]")
    nblk = (seq_len + block_size - 1) // block_size
    if len(block_table) < nblk:
        raise IndexError("list index out of range")      # what the reference's block_table[i // BS] raises
    bt = _as_block_table(block_table[:max(nblk, 1)] if len(block_table) else [0], query.device).view(1, -1)
    lens = torch.tensor([seq_len], dtype=torch.int32).to(query.device)
    return paged_attention_decode_batched(query.unsqueeze(0), kv_cache, bt, lens)[0]


# ---------------------------------------------------------------------------------------
# prefill (paged_attention.py:104-136)
# ---------------------------------------------------------------------------------------
def paged_attention_prefill_chunk(query: torch.Tensor, kv_cache: KvCacheArg, block_table: torch.Tensor,
                                  start_pos: int, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Causal attention of a query chunk ``[T, Hq, D]`` at positions ``[start_pos, start_pos + T)``
    over everything already in the cache for this sequence (history + the chunk itself, which the
    caller has written with ``reshape_and_cache``).  ``start_pos == 0`` is the reference prefill."""
    ca, cb = _split_cache(kv_cache)
    _require_cuda(query, "query")
    _require_cuda(block_table, "block_table")
    if query.dim() != 3:
        raise ValueError("query must be [T, Hq, D] in the cache dtype")
    if query.stride(2) != 1 and query.shape[2] > 1:
        query = query.contiguous()
    bt = block_table if block_table.dtype == torch.int32 and block_table.is_contiguous() else \
        block_table.to(torch.int32).contiguous()
    if out is None:
        out = torch.empty_like(query)
    t, hq, d = query.shape
    need = _lib.load().pa_prefill_workspace_bytes(t, hq, ca.shape[2] if ca.dim() >= 3 else 1, d)
    _lib.load_ext().prefill_chunk(query, ca, cb, bt, int(start_pos), out, _workspace(query.device, need))
    return out


def prefill_tensor_core_available(kv_cache: KvCacheArg, num_q_heads: int) -> bool:
    """True when prefill over this cache runs on the tcgen05 kernel (bf16, block_size 16,
    head_dim 64 / 128, g a power of two <= 16)."""
    ca, cb = _split_cache(kv_cache)
    return bool(_lib.load_ext().prefill_tc_available(ca, cb, int(num_q_heads)))


def paged_attention_prefill_varlen(query: torch.Tensor, kv_cache: KvCacheArg, block_tables: torch.Tensor,
                                   cu_q_lens: torch.Tensor, ctx_lens: torch.Tensor, max_q_len: int,
                                   out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """One launch for a batch of query chunks packed back to back (the engine's mixed
    chunked-prefill + decode step, mini_vllm.py:437-481): sequence ``s`` owns query tokens
    ``[cu_q_lens[s], cu_q_lens[s + 1])`` at positions ``[ctx_lens[s] - q_len, ctx_lens[s])`` and
    attends causally to ``[0, pos]`` through ``block_tables[s]``; its K / V (history and the chunk
    itself) must already be in the cache.  ``cu_q_lens`` int32 ``[B + 1]`` and ``ctx_lens`` int32
    ``[B]`` live on the device; ``max_q_len`` is a host int >= every chunk length.  Shapes outside
    the tensor-core path are looped per sequence through ``paged_attention_prefill_chunk``.
    The lengths are device tensors, so a block-table row too short for its context cannot raise the
    reference's IndexError here without a host sync: the kernel skips such a request (its rows of ``out``
    are left alone) and never reads the table past ``block_tables.shape[1]``."""
    ca, cb = _split_cache(kv_cache)
    _require_cuda(query, "query")
    if query.dim() == 3 and query.stride(2) != 1 and query.shape[2] > 1:
        query = query.contiguous()
    if out is None:
        out = torch.empty_like(query)
    rc = _lib.load_ext().prefill_varlen(query, ca, cb, block_tables, cu_q_lens, ctx_lens, int(max_q_len), out)
    if rc == -2:        # PA_ERR_UNSUPPORTED: not a tensor-core shape -> per-sequence CUDA path
        cu_h = cu_q_lens.tolist()
        cl_h = ctx_lens.tolist()
        for s in range(block_tables.shape[0]):
            q0, q1 = cu_h[s], cu_h[s + 1]
            if q1 > q0:
                paged_attention_prefill_chunk(query[q0:q1], kv_cache, block_tables[s], cl_h[s] - (q1 - q0),
                                              out=out[q0:q1])
    return out


def paged_attention_prefill(query: torch.Tensor, key: torch.Tensor, value: torch.Tensor, kv_cache: torch.Tensor,
                            block_table: List[int], block_size: int) -> torch.Tensor:
    """Reference signature (paged_attention.py:104-111): write K/V of positions 0..T-1 into the
    paged cache (in place), then causal GQA attention over the T tokens -> ``[T, Hq, D]``."""
    _require_cuda(query, "query")
    _require_cuda(kv_cache, "kv_cache")
    seq_len = query.shape[0]
    nblk = (seq_len + block_size - 1) // block_size
    if len(block_table) < nblk:
        raise IndexError("list index out of range")
    if kv_cache.dim() != 5 or kv_cache.shape[3] != block_size:
        raise ValueError("kv_cache must be [2, NB, Hkv, block_size, D]")
    if seq_len == 0:
        return torch.empty_like(query)
    bt = _as_block_table(block_table[:nblk], query.device)
    slots = compute_slot_mapping_device(bt, 0, seq_len, block_size)
    reshape_and_cache(key, value, kv_cache, slots)
    return paged_attention_prefill_chunk(query, kv_cache, bt, 0)


# ---------------------------------------------------------------------------------------
# block copy / swap (no reference code - definitional)
# ---------------------------------------------------------------------------------------
def _plane_groups(t: torch.Tensor, block_dim: int):
    """Split a pool tensor into groups of uniformly-strided planes of blocks.

    Returns (block_bytes, block_stride_bytes, [(byte_offset, num_planes, plane_stride_bytes), ...]).
    Everything after ``block_dim`` must be dense (one block = one contiguous run of bytes); the
    leading axes are collapsed into as few strided plane groups as their strides allow."""
    es = t.element_size()
    inner = 1
    for ax in range(t.dim() - 1, block_dim, -1):
        if t.shape[ax] != 1 and t.stride(ax) != inner:
            raise ValueError("the axes inside a block must be contiguous")
        inner *= t.shape[ax]
    block_bytes = inner * es
    block_stride = t.stride(block_dim) * es
    lead = [(t.shape[ax], t.stride(ax) * es) for ax in range(block_dim) if t.shape[ax] != 1]
    if not lead:
        return block_bytes, block_stride, [(0, 1, 0)]
    # collapse from the innermost leading axis outwards while strides nest exactly
    n, stride = lead[-1]
    k = len(lead) - 1
    while k > 0 and lead[k - 1][1] == stride * n:
        n *= lead[k - 1][0]
        k -= 1
    outer = lead[:k]
    groups = []
    for idx in itertools.product(*[range(s) for s, _ in outer]):
        off = sum(i * st for i, (_, st) in zip(idx, outer))
        groups.append((off, n, stride))
    return block_bytes, block_stride, groups


def _pairs_list(pairs) -> List[Tuple[int, int]]:
    if isinstance(pairs, torch.Tensor):
        pairs = pairs.detach().cpu().tolist()
    return [(int(s), int(d)) for s, d in pairs]


def copy_blocks(cache: Union[torch.Tensor, Sequence[torch.Tensor]], src_dst, block_dim: Optional[int] = None) -> None:
    """``cache[..., dst, ...] = cache[..., src, ...]`` on the block axis for every (src, dst) pair,
    over every plane (K/V, layers) of the pool, in one launch per plane group.  ``cache`` may be a
    pool ``[2, L, NB, ...]``, a layer ``[2, NB, ...]``, an engine cache ``[L, NB, ...]`` / ``[NB, ...]``
    or a sequence of such tensors; ``block_dim`` defaults to ``cache.dim() - 4``.  The destination set
    must not intersect the source set."""
    caches = list(cache) if isinstance(cache, (tuple, list)) else [cache]
    by_device = {}
    if isinstance(src_dst, torch.Tensor) and src_dst.is_cuda:
        # device-resident pairs: no host sync, and no disjointness check (caller's contract)
        dev_pairs = src_dst.to(torch.int64).contiguous()
        if dev_pairs.dim() != 2 or dev_pairs.shape[1] != 2:
            raise ValueError("src_dst must be [N, 2]")
        n = dev_pairs.shape[0]
        by_device[dev_pairs.device] = dev_pairs
        pl = None
    else:
        pl = _pairs_list(src_dst)
        n = len(pl)
        srcs = {s for s, _ in pl}
        if any(d in srcs and d != s for s, d in pl):
            raise ValueError("copy_blocks: destination blocks must not also be sources")
    if n == 0:
        return
    lib = _lib.load()
    launches = []           # (tensor for the device guard, base pointer, plane stride, planes, block stride, block bytes, pairs)
    for c in caches:
        _require_cuda(c, "cache")
        bd = c.dim() - 4 if block_dim is None else block_dim
        if pl is not None and any(not (0 <= s < c.shape[bd] and 0 <= d < c.shape[bd]) for s, d in pl):
            raise IndexError("copy_blocks: block id out of range")
        pairs_dev = by_device.get(c.device)
        if pairs_dev is None:
            if pl is None:
                pairs_dev = next(iter(by_device.values())).to(c.device)
            else:
                pairs_dev = torch.tensor(pl, dtype=torch.int64).to(c.device)
            by_device[c.device] = pairs_dev
        block_bytes, block_stride, groups = _plane_groups(c, bd)
        for off, planes, pstride in groups:
            launches.append([c, c.data_ptr() + off, pstride, planes, block_stride, block_bytes, pairs_dev])
    # Two single-plane groups of identical geometry (the engine's (k_cache, v_cache) pair, each [NB, BS, H, D]) are
    # two planes at a constant distance: ONE launch instead of two short ones (each pays its own ramp and tail).
    merged = []
    for ln in launches:
        prev = merged[-1] if merged else None
        if prev is not None and prev[3] == 1 and ln[3] == 1 and prev[4:6] == ln[4:6] and prev[6] is ln[6] \
                and prev[0].device == ln[0].device and ln[1] > prev[1] and (ln[1] - prev[1]) % 16 == 0:
            prev[2], prev[3] = ln[1] - prev[1], 2
        else:
            merged.append(ln)
    for c, base, pstride, planes, block_stride, block_bytes, pairs_dev in merged:
        with _DeviceOf(c):
            rc = lib.pa_copy_blocks(base, pstride, planes, block_stride, block_bytes, pairs_dev.data_ptr(), n, _stream_ptr())
            _lib.check(rc, "pa_copy_blocks")


class HostKVPool(BlockFifo):
    """Pinned host mirror of a device pool (same per-block layout, its own block count) - the
    swap space.  ``like`` is the device pool tensor; the block axis is ``block_dim``."""

    def __init__(self, like: torch.Tensor, num_blocks: int, block_dim: Optional[int] = None):
        bd = like.dim() - 4 if block_dim is None else block_dim
        shape = list(like.shape)
        shape[bd] = num_blocks
        super().__init__(num_blocks)
        self.num_blocks = num_blocks
        self.block_dim = bd
        self.kv_cache = torch.empty(shape, dtype=like.dtype, device='cpu', pin_memory=True)


_side_streams = {}


def _side_stream(device: torch.device) -> torch.cuda.Stream:
    s = _side_streams.get(device.index)
    if s is None:
        s = torch.cuda.Stream(device=device)
        _side_streams[device.index] = s
    return s


_SWAP_STAGING_BYTES = 256 << 20      # upper bound of the device staging area of a staged swap
_swap_stages = {}


def _swap_staging(device: torch.device, nbytes: int) -> torch.Tensor:
    """Per-device staging area of the staged swap (grown on demand, reused across calls; every
    use is ordered on the swap stream, so reuse needs no extra synchronisation)."""
    t = _swap_stages.get(device.index)
    if t is None or t.numel() < nbytes:
        t = torch.empty(nbytes, dtype=torch.uint8, device=device)
        _swap_stages[device.index] = t
    return t


def _swap(dst: torch.Tensor, src: torch.Tensor, pairs, block_dim: Optional[int], direction: int,
          stream: Optional[torch.cuda.Stream]) -> torch.cuda.Event:
    dev_t = src if direction == 0 else dst
    host_t = dst if direction == 0 else src
    _require_cuda(dev_t, "device cache")
    if host_t.is_cuda:
        raise ValueError("the host side of a swap must be a CPU (pinned) tensor")
    if not host_t.is_pinned():
        raise ValueError("the host pool must be pinned memory (HostKVPool allocates it that way)")
    bd = dev_t.dim() - 4 if block_dim is None else block_dim
    if dst.dtype != src.dtype or dst.dim() != src.dim() or \
            any(a != b for ax, (a, b) in enumerate(zip(dst.shape, src.shape)) if ax != bd):
        raise ValueError("swap needs pools of identical dtype and per-block layout")
    # pairs as one int64 [N, 2] array: bounds check, run counting and the C call all work on it
    # (a Python loop over 1024 pairs cost ~0.5 ms of host time before the first copy was issued)
    if isinstance(pairs, torch.Tensor):
        pairs = pairs.detach().cpu().numpy()
    pa = np.ascontiguousarray(np.asarray(pairs, dtype=np.int64).reshape(-1, 2))
    if pa.size and (pa.min() < 0 or pa[:, 0].max() >= src.shape[bd] or pa[:, 1].max() >= dst.shape[bd]):
        bad = pa[(pa[:, 0] < 0) | (pa[:, 0] >= src.shape[bd]) | (pa[:, 1] < 0) | (pa[:, 1] >= dst.shape[bd])][0]
        raise IndexError(f"swap pair ({int(bad[0])}, {int(bad[1])}) out of range")
    ev = torch.cuda.Event()
    side = stream or _side_stream(dev_t.device)
    with _DeviceOf(dev_t):
        side.wait_stream(torch.cuda.current_stream())      # KV writes issued so far are visible
        if len(pa):
            sb, sbs, sg = _plane_groups(src, bd)
            db, dbs, dg = _plane_groups(dst, bd)
            if sb != db or len(sg) != len(dg) or any(a[1] != b[1] for a, b in zip(sg, dg)):
                raise ValueError("swap: source and destination plane structures differ")
            lib = _lib.load()
            hcol = 1 if direction == 0 else 0
            step = np.diff(pa, axis=0) != 1
            both_runs = 1 + int(np.count_nonzero(step.any(axis=1)))
            host_runs = 1 + int(np.count_nonzero(step[:, hcol]))
            (h_bs, d_bs) = (dbs, sbs) if direction == 0 else (sbs, dbs)
            planes_total = sum(g[1] for g in sg)
            staged = (both_runs >= 8 and host_runs * planes_total < both_runs * len(sg) and h_bs == sb
                      and sb % 16 == 0 and d_bs % 16 == 0)
            if staged:
                # scattered device blocks: gather kernel -> dense staging -> one pinned copy per
                # (plane, run of host blocks); mirror image for swap-in.  Chunked by staging size.
                per_pair = sb * max(g[1] for g in sg)
                chunk = max(1, min(len(pa), _SWAP_STAGING_BYTES // per_pair, 65535))
                stage = _swap_staging(dev_t.device, chunk * per_pair)
                for at in range(0, len(pa), chunk):
                    part = pa[at:at + chunk]
                    with torch.cuda.stream(side):
                        # pageable -> device: staged by the driver before the call returns and
                        # ordered on the swap stream, so neither buffer outlives its use
                        pairs_dev = torch.from_numpy(part).to(dev_t.device)
                    for (soff, planes, sps), (doff, _, dps) in zip(sg, dg):
                        (d_ptr, d_ps, h_ptr, h_ps) = (src.data_ptr() + soff, sps, dst.data_ptr() + doff, dps) \
                            if direction == 0 else (dst.data_ptr() + doff, dps, src.data_ptr() + soff, sps)
                        rc = lib.pa_swap_blocks_staged(d_ptr, d_ps, d_bs, h_ptr, h_ps, h_bs, planes, sb,
                                                       part.ctypes.data, pairs_dev.data_ptr(), len(part), direction,
                                                       stage.data_ptr(), stage.numel(), side.cuda_stream)
                        _lib.check(rc, "pa_swap_blocks_staged")
            else:
                for (soff, planes, sps), (doff, _, dps) in zip(sg, dg):
                    rc = lib.pa_swap_blocks(dst.data_ptr() + doff, dps, dbs, src.data_ptr() + soff, sps, sbs,
                                            planes, sb, pa.ctypes.data, len(pa), direction, side.cuda_stream)
                    _lib.check(rc, "pa_swap_blocks")
        ev.record(side)
    return ev


def swap_out(dev_cache: torch.Tensor, host_cache: torch.Tensor, src_dst, block_dim: Optional[int] = None,
             stream: Optional[torch.cuda.Stream] = None) -> torch.cuda.Event:
    """Device -> pinned host: ``host_cache[..., dst, ...] = dev_cache[..., src, ...]`` with
    ``cudaMemcpyAsync`` on a side stream.  Returns the event recorded after the copies; wait on it
    (``torch.cuda.current_stream().wait_event(ev)`` or ``ev.synchronize()``) before reusing the
    device blocks or reading the host copy."""
    return _swap(host_cache, dev_cache, src_dst, block_dim, 0, stream)


def swap_in(host_cache: torch.Tensor, dev_cache: torch.Tensor, src_dst, block_dim: Optional[int] = None,
            stream: Optional[torch.cuda.Stream] = None) -> torch.cuda.Event:
    """Pinned host -> device: ``dev_cache[..., dst, ...] = host_cache[..., src, ...]``.  The compute
    stream must wait on the returned event before attending over the swapped-in blocks."""
    return _swap(dev_cache, host_cache, src_dst, block_dim, 1, stream)
