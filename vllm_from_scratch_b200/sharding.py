"""Multi-GPU partitioning of the decode path.

One process per GPU; every rank owns its block pool, allocator and block tables.  Decode
attention is independent per (sequence, KV head), so there are two ways to cut it:

* by SEQUENCE (default): rank r runs sequences ``shard_sequences(B, world, r)``; block ids are
  rank-local; no collective on the data path (an optional all-gather only if one consumer needs
  the whole batch);
* by KV HEAD (long context / small batch, world <= Hkv): rank r holds KV heads
  ``shard_kv_heads(Hkv, world, r)`` of EVERY block and the matching query heads; all ranks share
  the block tables.  The only collective is one all-gather of the ``[B, Hq / world, D]`` outputs.
  ``HeadGather`` owns the gather buffer ``[world, B, Hq / world, D]``: the decode kernel writes its
  slice straight into ``local_out()`` (a contiguous slot of that buffer - the C ABI takes output
  strides, so no staging copy) and ``all_gather()`` completes the buffer in place with ONE
  ``all_gather_into_tensor`` (NCCL over NVLink on GPUs, gloo in the CPU tests).

This module is host-side arithmetic plus the one ``torch.distributed`` call; it never touches
the kernels and works on CPU tensors too (that is how the gloo tests cover it).
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
import torch.distributed as dist


def shard_sequences(batch: int, world: int, rank: int) -> range:
    """Contiguous, balanced split of ``range(batch)``: the first ``batch % world`` ranks get one more."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world {world}")
    base, extra = divmod(batch, world)
    lo = rank * base + min(rank, extra)
    return range(lo, lo + base + (1 if rank < extra else 0))


def shard_kv_heads(num_kv_heads: int, world: int, rank: int) -> range:
    """KV heads of rank ``rank``; ``world`` must divide ``num_kv_heads``."""
    if num_kv_heads % world:
        raise ValueError(f"KV-head sharding needs world ({world}) to divide num_kv_heads ({num_kv_heads})")
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world {world}")
    per = num_kv_heads // world
    return range(rank * per, (rank + 1) * per)


def shard_q_heads(num_q_heads: int, num_kv_heads: int, world: int, rank: int) -> range:
    """Query heads that belong to rank ``rank``'s KV heads (GQA groups stay together)."""
    g = num_q_heads // num_kv_heads
    kv = shard_kv_heads(num_kv_heads, world, rank)
    return range(kv.start * g, kv.stop * g)


def slice_kv_heads(kv_cache: torch.Tensor, num_kv_heads: int, world: int, rank: int, head_dim_axis: int) -> torch.Tensor:
    """The rank's head slice of a full cache tensor (used to build per-rank pools from a global
    one in tests and benches; production ranks allocate only their slice)."""
    kv = shard_kv_heads(num_kv_heads, world, rank)
    return kv_cache.narrow(head_dim_axis, kv.start, len(kv)).contiguous()


class HeadGather:
    """Gather buffer for KV-head-sharded decode outputs."""

    def __init__(self, batch: int, num_q_heads: int, head_dim: int, dtype, device, world: Optional[int] = None,
                 rank: Optional[int] = None, group=None):
        self.group = group
        self.world = world if world is not None else dist.get_world_size(group)
        self.rank = rank if rank is not None else dist.get_rank(group)
        if num_q_heads % self.world:
            raise ValueError("world must divide num_q_heads")
        self.batch, self.num_q_heads, self.head_dim = batch, num_q_heads, head_dim
        self.buf = torch.empty(self.world, batch, num_q_heads // self.world, head_dim, dtype=dtype, device=device)

    def local_out(self) -> torch.Tensor:
        """``[B, Hq / world, D]`` - pass as ``out=`` to ``paged_attention_decode_batched``."""
        return self.buf[self.rank]

    def all_gather(self, async_op: bool = False):
        """Complete the buffer in place; returns the work handle when ``async_op``."""
        if self.world == 1:
            return None
        return dist.all_gather_into_tensor(self.buf.view(-1), self.buf[self.rank].reshape(-1), group=self.group,
                                           async_op=async_op)

    def full_view(self) -> torch.Tensor:
        """``[B, world, Hq / world, D]`` view of the gathered outputs, no copy: ``[b, r, j]`` is query
        head ``r * Hq / world + j`` of sequence ``b``."""
        return self.buf.permute(1, 0, 2, 3)

    def full(self) -> torch.Tensor:
        """``[B, Hq, D]`` with heads in global order (one small copy of B * Hq * D elements when
        world > 1: NCCL needs contiguous per-rank chunks, so the head axis cannot be innermost)."""
        return self.full_view().reshape(self.batch, self.num_q_heads, self.head_dim)


def gather_sequence_outputs(local_out: torch.Tensor, batch: int, world: int, rank: int, group=None) -> torch.Tensor:
    """Optional, sequence-sharded mode: all ranks' ``[B_r, Hq, D]`` -> ``[B, Hq, D]`` (ragged shards
    are padded to the largest).  Not on the attention data path."""
    if world == 1:
        return local_out
    per = len(shard_sequences(batch, world, 0))
    pad = torch.zeros((per,) + tuple(local_out.shape[1:]), dtype=local_out.dtype, device=local_out.device)
    pad[:local_out.shape[0]] = local_out
    buf = torch.empty((world, per) + tuple(local_out.shape[1:]), dtype=local_out.dtype, device=local_out.device)
    dist.all_gather_into_tensor(buf.view(-1), pad.view(-1), group=group)
    parts = [buf[r, :len(shard_sequences(batch, world, r))] for r in range(world)]
    return torch.cat(parts, dim=0)
