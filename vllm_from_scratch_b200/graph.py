"""One decode step of the hot path as a CUDA graph.

The engine's per-step work around the kernels - building tensors from Python lists, four to six
small H2D copies, two launches, one D2H copy (05_mini_vllm/mini_vllm.py:729-748 plus the
attention call site) - costs more host time than the 90 us the decode kernel runs.
``DecodeStep`` fixes the shapes once, keeps ONE pinned staging buffer for everything the step
reads from the host (q, the new K / V rows, their slots, block tables, sequence lengths) and ONE
pinned result buffer, and captures

    H2D(staging) -> reshape_and_cache -> paged_attention_decode -> D2H(out)

into a ``torch.cuda.CUDAGraph``.  With ``zero_copy`` (default) the two largest per-step operands
skip the staging copies altogether: the decode kernel fetches ``q`` from pinned host memory with
its own bulk copies (a unit's Q rows are requested ~8 us before they are needed, so the PCIe
latency hides behind the KV stream) and writes ``out`` straight into the pinned result buffer.
A step is then: fill the pinned views, ``replay()``, ``wait()``.  The block tables stay device-resident between steps if the caller only rewrites
the entries that changed in the pinned view (the whole staging buffer is one copy either way).

Everything here is plumbing (memory, streams, graph capture); the work is done by the same C-ABI
kernels as the eager path.
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import torch

from . import paged_attention as P


class DecodeStep:
    def __init__(self, kv_cache: P.KvCacheArg, batch: int, num_q_heads: int, max_blocks: int,
                 write_kv: bool = True, out: Optional[torch.Tensor] = None, zero_copy: bool = True):
        """``zero_copy``: the decode kernel reads ``q`` from and writes ``out`` to pinned host memory
        in place (no staging copy for the two largest per-step operands); everything else is staged.
        A caller-provided device ``out`` (e.g. a slot of a gather buffer) disables it for the output."""
        desc, kplane = P._describe_cache(kv_cache)
        self.kv_cache = kv_cache
        self.device = kplane.device
        self.dtype = kplane.dtype
        self.batch, self.num_q_heads, self.max_blocks = batch, num_q_heads, max_blocks
        self.write_kv = write_kv
        hkv, d = desc.num_kv_heads, desc.head_dim
        es = kplane.element_size()
        self.zero_copy = zero_copy
        fields = [] if zero_copy else [("q", (batch, num_q_heads, d), self.dtype, es)]
        if write_kv:
            fields += [("k", (batch, hkv, d), self.dtype, es), ("v", (batch, hkv, d), self.dtype, es),
                       ("slots", (batch,), torch.int64, 8)]
        fields += [("block_tables", (batch, max_blocks), torch.int32, 4), ("seq_lens", (batch,), torch.int32, 4)]
        offs: Dict[str, Tuple[int, int]] = {}
        at = 0
        for name, shape, _, esize in fields:
            at = (at + 255) // 256 * 256
            n = esize
            for s in shape:
                n *= s
            offs[name] = (at, n)
            at += n
        self.h2d_bytes = at
        self._stage_h = torch.empty(at, dtype=torch.uint8).pin_memory()
        self._stage_d = torch.empty(at, dtype=torch.uint8, device=self.device)
        self.host: Dict[str, torch.Tensor] = {}
        self._dev: Dict[str, torch.Tensor] = {}
        for name, shape, dtype, _ in fields:
            lo, n = offs[name]
            self.host[name] = self._stage_h[lo:lo + n].view(dtype).view(shape)
            self._dev[name] = self._stage_d[lo:lo + n].view(dtype).view(shape)
        self.host["block_tables"].zero_()
        self.host["seq_lens"].zero_()
        if write_kv:
            self.host["slots"].fill_(-1)
        self.out = torch.empty((batch, num_q_heads, d), dtype=self.dtype).pin_memory()
        self._out_in_place = zero_copy and out is None
        if self._out_in_place:
            self.out_dev = self.out
        else:
            self.out_dev = out if out is not None else torch.empty((batch, num_q_heads, d), dtype=self.dtype,
                                                                   device=self.device)
        if zero_copy:
            self.host["q"] = torch.empty((batch, num_q_heads, d), dtype=self.dtype).pin_memory()
            self._dev["q"] = self.host["q"]
            self.h2d_bytes += self.host["q"].numel() * es            # still crosses PCIe every step
        self.d2h_bytes = self.out.numel() * es
        self._ws = P.new_decode_workspace(batch, num_q_heads, hkv, d, self.device)
        self._graph: Optional[torch.cuda.CUDAGraph] = None
        self._done = torch.cuda.Event()

    # ---- the step, eager ----------------------------------------------------------------
    def _enqueue(self) -> None:
        self._stage_d.copy_(self._stage_h, non_blocking=True)
        dv = self._dev
        if self.write_kv:
            P.reshape_and_cache(dv["k"], dv["v"], self.kv_cache, dv["slots"])
        P.paged_attention_decode_batched(dv["q"], self.kv_cache, dv["block_tables"], dv["seq_lens"],
                                         out=self.out_dev, workspace=self._ws)
        if not self._out_in_place:
            self.out.copy_(self.out_dev, non_blocking=True)

    def capture(self) -> "DecodeStep":
        """Warm the step up once (tensor maps, function attributes) and capture it."""
        with torch.cuda.device(self.device):
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                self._enqueue()
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._enqueue()
            self._graph = g
        return self

    def replay(self) -> None:
        """Launch the step on the current stream (one graph launch when captured)."""
        if self._graph is not None:
            self._graph.replay()
        else:
            self._enqueue()
        self._done.record()

    def wait(self) -> torch.Tensor:
        """Block until the step's result is in ``self.out`` (pinned host memory) and return it."""
        self._done.synchronize()
        return self.out
