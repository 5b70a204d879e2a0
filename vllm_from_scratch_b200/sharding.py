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


class _DevMem:
    """Raw device memory as a torch tensor (``__cuda_array_interface__``), no copy, no ownership."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3}


class PeerGather:
    """KV-head-sharded decode with the output gather FUSED into the decode kernel.

    ``HeadGather`` finishes the step with an NCCL all-gather: one more launch and a collective whose
    latency (tens of microseconds through the host launch path) exceeds the kernel's at 8 GPUs.
    Here every rank's gather buffer lives in peer-visible memory (``pa_peer_alloc`` + CUDA IPC, one
    process per GPU); the decode kernel stores each finished output row into its own slot of EVERY
    rank's buffer over NVLink and bumps its arrival counter there (``pa_decode_gather``); one tiny
    ``pa_peer_wait`` kernel on the consumer's stream makes the gathered ``[world, B, Hq / world, D]``
    visible.  Two buffer sets alternate (a fast rank may be one step ahead of a slow one; see
    include/pa_b200.h).  Default ``protocol='ll'``: every 4-byte output word travels as {data, step
    number} in one 8-byte store into an LL area of the receiver, the sender fences and counts nothing,
    and the receiver's ``pa_peer_unpack`` kernel polls the step numbers and fills the gather set;
    ``'counter'`` is the plain-stores + ``fence.sys`` + arrival-counter form.  ``torch.distributed`` is used only to exchange the 64-byte IPC handles.
    """

    def __init__(self, batch: int, num_q_heads: int, num_kv_heads: int, head_dim: int, device, world: Optional[int] = None,
                 rank: Optional[int] = None, group=None, timeout_us: int = 2_000_000, protocol: str = "ll"):
        import ctypes

        from . import _lib
        self.group = group
        self.world = world if world is not None else dist.get_world_size(group)
        self.rank = rank if rank is not None else dist.get_rank(group)
        if not (1 <= self.world <= 8):
            raise ValueError("PeerGather supports 1..8 ranks of one node")
        if num_kv_heads % self.world or num_q_heads % num_kv_heads:
            raise ValueError("world must divide num_kv_heads, num_kv_heads must divide num_q_heads")
        self.batch, self.num_q_heads, self.head_dim = batch, num_q_heads, head_dim
        self.hq_local = num_q_heads // self.world
        self.segments = batch * (num_kv_heads // self.world)          # arrivals per rank per step
        self.device = torch.device(device)
        self.timeout_us = timeout_us
        self.step = 0
        self._lib = _lib.load()
        slot = batch * self.hq_local * head_dim * 2                    # bf16
        if protocol not in ("ll", "counter"):
            raise ValueError("protocol must be 'll' (flag travels with every data word, nothing fenced on the sender) "
                             "or 'counter' (plain stores + fence + arrival counters)")
        self.protocol = protocol
        self._slot_words = slot // 4
        self._set_bytes = self.world * slot
        self._ll_off = (2 * self._set_bytes + 255) // 256 * 256         # LL areas: 8 bytes per 4-byte output word
        self._ll_set_bytes = 2 * self._set_bytes
        self._sig_off = (self._ll_off + 2 * self._ll_set_bytes + 255) // 256 * 256
        total = self._sig_off + 256
        ptr, handle = ctypes.c_void_p(), ctypes.create_string_buffer(64)
        with torch.cuda.device(self.device):
            _lib.check(self._lib.pa_peer_alloc(total, ctypes.byref(ptr), handle), "pa_peer_alloc")
        self._base = ptr.value
        handles = [None] * self.world
        if self.world > 1:
            dist.all_gather_object(handles, handle.raw, group=group)
        self._remote = {}
        for r in range(self.world):
            if r == self.rank:
                continue
            rp = ctypes.c_void_p()
            with torch.cuda.device(self.device):
                _lib.check(self._lib.pa_peer_open(ctypes.create_string_buffer(handles[r], 64), ctypes.byref(rp)), "pa_peer_open")
            self._remote[r] = rp.value
        mem = torch.as_tensor(_DevMem(self._base, total), device=self.device)
        self.sets = [mem[i * self._set_bytes:(i + 1) * self._set_bytes].view(torch.bfloat16)
                     .view(self.world, batch, self.hq_local, head_dim) for i in range(2)]
        self.sig = mem[self._sig_off:self._sig_off + 4 * self.world].view(torch.int32)
        self.status = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.epoch = torch.zeros(1, dtype=torch.int32, device=self.device)     # steps waited for, kept on the device
        self._done = torch.zeros(1, dtype=torch.int32, device=self.device)     # block counter of the unpack kernel
        self._peers = []
        for i in range(2):
            pg = _lib.PaPeerGather()
            pg.num_remote = self.world - 1
            my_slot = i * self._set_bytes + self.rank * slot
            for j, (r, base) in enumerate(sorted(self._remote.items())):
                if protocol == "ll":
                    pg.remote_out[j] = base + self._ll_off + i * self._ll_set_bytes + self.rank * slot * 2
                else:
                    pg.remote_out[j] = base + my_slot
                    pg.remote_sig[j] = base + self._sig_off + 4 * self.rank
            if protocol == "ll":
                pg.ll_epoch = self.epoch.data_ptr()
            else:
                pg.local_sig = self._base + self._sig_off + 4 * self.rank
            self._peers.append(pg)
        if self.world > 1:
            dist.barrier(group=group)                                   # every rank has mapped every buffer

    def local_out(self) -> torch.Tensor:
        """``[B, Hq / world, D]``: this rank's slot of the buffer set the NEXT ``decode`` fills."""
        return self.sets[self.step % 2][self.rank]

    def decode(self, query: torch.Tensor, kv_cache, block_tables: torch.Tensor, seq_lens: torch.Tensor,
               workspace: Optional[torch.Tensor] = None, parity: Optional[int] = None) -> torch.Tensor:
        """One head-sharded decode step on the current stream; returns the gathered outputs
        ``[world, B, Hq / world, D]`` (valid for later work on this stream; use before the step after
        next overwrites the set).  Both launches are graph-capturable (the step number the wait needs
        lives on the device): capture one graph per ``parity`` (0, 1) and replay them ALTERNATELY,
        starting with 0 - all ranks must run the same sequence of steps."""
        from . import _lib
        from . import paged_attention as P
        i = self.step % 2 if parity is None else parity
        P.paged_attention_decode_batched(query, kv_cache, block_tables, seq_lens, out=self.sets[i][self.rank],
                                         workspace=workspace, peers=self._peers[i])
        if parity is None:
            self.step += 1
        with torch.cuda.device(self.device):
            st = torch.cuda.current_stream().cuda_stream
            if self.protocol == "ll":
                rc = self._lib.pa_peer_unpack(self._base + self._ll_off + i * self._ll_set_bytes, self.sets[i].data_ptr(),
                                              self.world, self.rank, self._slot_words, self.epoch.data_ptr(),
                                              self._done.data_ptr(), self.timeout_us, self.status.data_ptr(), st)
            else:
                rc = self._lib.pa_peer_wait_next(self.sig.data_ptr(), self.world, self.epoch.data_ptr(), self.segments,
                                                 self.timeout_us, self.status.data_ptr(), st)
        _lib.check(rc, "pa_peer_unpack" if self.protocol == "ll" else "pa_peer_wait_next")
        return self.sets[i]

    def check(self) -> None:
        """Synchronises and raises if a wait timed out (a peer never delivered its rows)."""
        late = int(self.status.item())
        if late:
            raise RuntimeError(f"PeerGather: rank {late - 1} did not deliver its outputs within {self.timeout_us} us")

    def full(self, gathered: torch.Tensor) -> torch.Tensor:
        """``[B, Hq, D]`` with heads in global order (one small copy)."""
        return gathered.permute(1, 0, 2, 3).reshape(self.batch, self.num_q_heads, self.head_dim)

    def close(self) -> None:
        if self._base is None:
            return
        torch.cuda.synchronize(self.device)
        if self.world > 1:
            dist.barrier(group=self.group)                              # nobody is still writing into a peer
        self.sets, self.sig = [], None
        with torch.cuda.device(self.device):
            for base in self._remote.values():
                self._lib.pa_peer_close(base)
            self._lib.pa_peer_free(self._base)
        self._remote, self._base = {}, None
