"""Replay side of the engine-loop golden trace (test infrastructure, not product code).

``tests/golden/make_engine_golden.py`` runs the UNMODIFIED reference engine
(``05_mini_vllm/mini_vllm.py:641-783``: ``MiniVLLM`` = ``Scheduler`` + ``BlockAllocator`` +
``MiniTransformer`` + ``Sampler``) in the build container with ``engine_adapter`` + ``kv_manager``
installed over the CPU oracle, and records every step the real scheduler produced (tokens, positions,
slot mappings, block tables, prefill flags, swap traffic) together with the real model's logits.
``/root/reference`` does not exist on the GPU box, so there the recorded steps are replayed:

* ``ReplayScheduler`` hands the recorded ``ScheduledRequest``s to the SAME ``engine_adapter.install``
  wrapper the real engine uses (and first pushes the recorded swap-out / swap-in pairs through the
  CUDA swap path);
* ``ReplicaTransformer`` is a small decoder with the reference model's state_dict names and
  arithmetic (RMSNorm ``w * x / sqrt(mean(x^2) + eps)`` :344-352, rotary tables :360-374, SwiGLU
  :490-499, pre-norm residual layers :502-514, last-token logits :573-585), written from that
  description; its attention forward is absent - it only works with the adapter installed.  Weights
  come from ``engine_cases.det_weights`` on both sides (sha256 pinned in the golden file).
* ``ReplayEngine.step`` builds the model inputs exactly like ``MiniVLLM.step`` (:721-749).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F


class RMSNorm(nn.Module):
    def __init__(self, dim: int, eps: float = 1e-6):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(dim))
        self.eps = eps

    def forward(self, x):
        return self.weight * (x / x.pow(2).mean(-1, keepdim=True).add(self.eps).sqrt())


class Rotary(nn.Module):
    def __init__(self, dim: int, max_len: int = 8192, base: float = 10000.0):
        super().__init__()
        inv_freq = 1.0 / (base ** (torch.arange(0, dim, 2).float() / dim))
        freqs = torch.outer(torch.arange(max_len).float(), inv_freq)
        emb = torch.cat([freqs, freqs], dim=-1)
        self.register_buffer("inv_freq", inv_freq)
        self.register_buffer("cos_cached", emb.cos())
        self.register_buffer("sin_cached", emb.sin())

    def forward(self, positions):
        return self.cos_cached[positions], self.sin_cached[positions]


class Attn(nn.Module):
    def __init__(self, hidden: int, heads: int):
        super().__init__()
        self.num_heads, self.head_dim = heads, hidden // heads
        self.qkv = nn.Linear(hidden, 3 * hidden, bias=False)
        self.out_proj = nn.Linear(hidden, hidden, bias=False)
        self.rope = Rotary(self.head_dim)

    def forward(self, *a, **k):
        raise RuntimeError("ReplicaTransformer needs engine_adapter.install(engine)")


class SwiGLU(nn.Module):
    def __init__(self, hidden: int, inter: int):
        super().__init__()
        self.gate = nn.Linear(hidden, inter, bias=False)
        self.up = nn.Linear(hidden, inter, bias=False)
        self.down = nn.Linear(inter, hidden, bias=False)

    def forward(self, x):
        return self.down(F.silu(self.gate(x)) * self.up(x))


class Layer(nn.Module):
    def __init__(self, hidden: int, heads: int, inter: int):
        super().__init__()
        self.attn = Attn(hidden, heads)
        self.ffn = SwiGLU(hidden, inter)
        self.norm1, self.norm2 = RMSNorm(hidden), RMSNorm(hidden)

    def forward(self, x, positions, kv_cache, slot_mapping, seq_lens, is_prefill):
        x = x + self.attn(self.norm1(x), positions, kv_cache, slot_mapping, seq_lens, is_prefill)
        return x + self.ffn(self.norm2(x))


class ReplicaTransformer(nn.Module):
    def __init__(self, vocab_size, hidden_size, num_layers, num_heads, intermediate_size, num_kv_blocks, block_size):
        super().__init__()
        self.embedding = nn.Embedding(vocab_size, hidden_size)
        self.layers = nn.ModuleList([Layer(hidden_size, num_heads, intermediate_size) for _ in range(num_layers)])
        self.norm = RMSNorm(hidden_size)
        self.lm_head = nn.Linear(hidden_size, vocab_size, bias=False)
        hd = hidden_size // num_heads
        self.register_buffer("k_cache", torch.zeros(num_layers, num_kv_blocks, block_size, num_heads, hd))
        self.register_buffer("v_cache", torch.zeros(num_layers, num_kv_blocks, block_size, num_heads, hd))

    def forward(self, token_ids, positions, slot_mapping, seq_lens, is_prefill):
        x = self.embedding(token_ids)
        for i, layer in enumerate(self.layers):
            x = layer(x, positions, (self.k_cache[i], self.v_cache[i]), slot_mapping, seq_lens, is_prefill)
        x = self.norm(x)
        last, off = [], 0
        for n in seq_lens:
            last.append(off + n - 1)
            off += n
        return self.lm_head(x[last])


# ---------------------------------------------------------------------------------------
# the recorded trace
# ---------------------------------------------------------------------------------------
@dataclass
class _Req:
    request_id: str
    block_table: List[int]


@dataclass
class _Sched:
    request: _Req
    tokens: List[int]
    positions: List[int]
    is_prefill: bool
    slot_mapping: List[int]


class Trace:
    """Reader of tests/golden/engine_<case>.npz (layout written by make_engine_golden.py)."""

    def __init__(self, path: str):
        z = np.load(path, allow_pickle=False)
        self.z = z
        self.num_steps = len(z["step_ptr"]) - 1
        self.weights_sha = str(z["weights_sha"])
        self.full_steps = {int(s): i for i, s in enumerate(z["full_steps"])}

    def step(self, s: int) -> List[_Sched]:
        z = self.z
        out = []
        for r in range(int(z["step_ptr"][s]), int(z["step_ptr"][s + 1])):
            t0, t1 = int(z["tok_ptr"][r]), int(z["tok_ptr"][r + 1])
            b0, b1 = int(z["bt_ptr"][r]), int(z["bt_ptr"][r + 1])
            start = int(z["start_pos"][r])
            out.append(_Sched(_Req(f"req-{int(z['req_ids'][r])}", [int(b) for b in z["bt_flat"][b0:b1]]),
                              [int(t) for t in z["tokens_flat"][t0:t1]], list(range(start, start + t1 - t0)),
                              bool(z["is_prefill"][r]), [int(x) for x in z["slots_flat"][t0:t1]]))
        return out

    def events(self, s: int):
        """[(kind 'out' | 'in', cache 'k' | 'v', [(src, dst), ...], poisoned device blocks)] applied
        before step s's schedule."""
        z = self.z
        out = []
        for e in np.nonzero(z["ev_step"] == s)[0]:
            p0, p1 = int(z["ev_ptr"][e]), int(z["ev_ptr"][e + 1])
            d0, d1 = int(z["ev_dead_ptr"][e]), int(z["ev_dead_ptr"][e + 1])
            out.append(("in" if z["ev_kind"][e] else "out", "v" if z["ev_cache"][e] else "k",
                        [(int(a), int(b)) for a, b in z["ev_pairs"][p0:p1]], [int(b) for b in z["ev_dead"][d0:d1]]))
        return out

    def expected(self, s: int):
        """(next_tokens [R], top8 values [R, 8], top8 indices [R, 8], full logits [R, V] or None)."""
        z = self.z
        r0, r1 = int(z["step_ptr"][s]), int(z["step_ptr"][s + 1])
        full = None
        if s in self.full_steps:
            i = self.full_steps[s]
            full = z["full_logits"][int(z["full_ptr"][i]):int(z["full_ptr"][i + 1])]
        return z["next_tok"][r0:r1], z["top_val"][r0:r1], z["top_idx"][r0:r1], full


class ReplayScheduler:
    def __init__(self, engine, trace: Trace):
        self.engine, self.trace, self.cursor = engine, trace, 0

    def schedule(self):
        s = self.cursor
        self.engine.apply_events(self.trace.events(s))
        self.cursor += 1
        return self.trace.step(s)

    @property
    def has_unfinished(self):
        return self.cursor < self.trace.num_steps


class ReplayEngine:
    """``MiniVLLM``-shaped object (``device``, ``model``, ``scheduler``) for ``engine_adapter.install``."""

    def __init__(self, cfg, weights, trace: Trace, device, dtype, ops):
        self.device = torch.device(device)
        model = ReplicaTransformer(**{k: cfg[k] for k in ("vocab_size", "hidden_size", "num_layers", "num_heads",
                                                          "intermediate_size", "num_kv_blocks", "block_size")})
        missing = model.load_state_dict(weights, strict=False)
        assert not missing.unexpected_keys and all(k.endswith(("_cache", "_cached", "inv_freq")) for k in missing.missing_keys)
        self.model = model.to(dtype).to(self.device)           # cast on the CPU first: same rounding as the golden side
        self.scheduler = ReplayScheduler(self, trace)
        self.ops = ops
        self.host = None
        self.num_host_blocks = cfg["num_host_blocks"]
        self.swapped_blocks = 0

    def apply_events(self, events):
        if not events:
            return
        P = self.ops
        m = self.model
        if self.host is None:
            self.host = {"k": P.HostKVPool(m.k_cache, self.num_host_blocks, block_dim=1),
                         "v": P.HostKVPool(m.v_cache, self.num_host_blocks, block_dim=1)}
        for kind, which, pairs, dead in events:
            dev = m.k_cache if which == "k" else m.v_cache
            host = self.host[which].kv_cache
            ev = P.swap_out(dev, host, pairs, block_dim=1) if kind == "out" else P.swap_in(host, dev, pairs, block_dim=1)
            if ev is not None:
                ev.wait()
            if dead:
                # the recorded run poisoned these swapped-out device blocks too (make_engine_golden.py:
                # the ones that are not prefix-cache candidates): a swap path that silently did nothing
                # cannot reproduce its logits
                dev[:, dead] = float("nan")
            self.swapped_blocks += len(pairs)

    @torch.no_grad()
    def step(self):
        scheduled = self.scheduler.schedule()                   # adapter-wrapped: builds the step's metadata
        toks = [t for s in scheduled for t in s.tokens]
        pos = [p for s in scheduled for p in s.positions]
        slots = [x for s in scheduled for x in s.slot_mapping]
        dev = self.device
        return self.model(torch.tensor(toks, dtype=torch.long, device=dev), torch.tensor(pos, dtype=torch.long, device=dev),
                          torch.tensor(slots, dtype=torch.long, device=dev), [len(s.tokens) for s in scheduled],
                          [s.is_prefill for s in scheduled])
