"""Experiment (not part of the product): where a prefill work tile's time goes.  Per-work-tile globaltimer
stamps of pa_prefill_tc2_kernel (pa_prefill_debug_times) on the bench_prefill shapes.  The kernel is persistent
(one CTA per SM walks the work tiles), so "CTA" below reads "work tile": [1] = warpgroup A starts on the tile (its
previous tile's output is stored), [3] = its Q tile staged (BEFORE [1] for every tile but a CTA's first: it is
staged under the previous tile's last softmax pass), [4] first S ready, [5] O complete, [6] outputs stored; the
same-SM gap is the time between two consecutive tiles of one CTA."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import bench_prefill as BP  # noqa: E402
from vllm_from_scratch_b200 import _lib, paged_attention as P  # noqa: E402

lib = _lib.load()
HQ, HKV, D, BS = BP.HQ, BP.HKV, BP.D, BP.BS
dev = "cuda"
for name in (sys.argv[1:] or ["batch2k", "full8k"]):
    specs = BP.SHAPES[name]
    g = torch.Generator(device=dev).manual_seed(0)
    nblk = [(c + BS - 1) // BS for c, _ in specs]
    nb = sum(nblk) + 16
    kv = torch.randn(2, nb, HKV, BS, D, generator=g, device=dev, dtype=torch.float32).to(torch.bfloat16)
    perm = torch.randperm(nb, generator=torch.Generator().manual_seed(0)).to(torch.int32)
    bts = torch.zeros(len(specs), max(nblk), dtype=torch.int32)
    at = 0
    for i, n in enumerate(nblk):
        bts[i, :n] = perm[at:at + n]
        at += n
    cu = torch.tensor([0] + list(torch.tensor([q for _, q in specs]).cumsum(0)), dtype=torch.int32)
    ctx = torch.tensor([c for c, _ in specs], dtype=torch.int32)
    q = torch.randn(int(cu[-1]), HQ, D, generator=g, device=dev, dtype=torch.float32).to(torch.bfloat16)
    out = torch.empty_like(q)
    maxq = max(ql for _, ql in specs)
    args = (q, kv, bts.to(dev), cu.to(dev), ctx.to(dev), maxq)
    for _ in range(3):
        P.paged_attention_prefill_varlen(*args, out=out)
    torch.cuda.synchronize()
    gx = (maxq + 63) // 64                                     # 2 * R = 64 tokens per CTA at g = 4
    ncta = gx * HKV * len(specs)
    buf = torch.zeros(ncta * 8, dtype=torch.int64, device=dev)
    lib.pa_prefill_debug_times(buf.data_ptr())
    P.paged_attention_prefill_varlen(*args, out=out)
    torch.cuda.synchronize()
    lib.pa_prefill_debug_times(None)
    r = buf.cpu().numpy().reshape(ncta, 8)
    r = r[r[:, 1] > 0]
    t0 = r[:, 1].min()
    us = lambda a: a / 1e3                                     # noqa: E731
    pro, qst, first, body, epi = (us(r[:, 2] - r[:, 1]), us(r[:, 3] - r[:, 2]), us(r[:, 4] - r[:, 1]),
                                  us(r[:, 5] - r[:, 4]), us(r[:, 6] - r[:, 5]))
    tiles = r[:, 7]
    total = us(r[:, 6] - r[:, 1])
    print(f"== {name}: {len(r)} CTAs, kernel span {us(r[:, 6].max() - t0):.1f} us, tiles/CTA mean {tiles.mean():.1f}")
    print(f"   entry->prologue {np.median(pro):.2f} us | Q staged +{np.median(qst):.2f} | entry->first S {np.median(first):.2f} "
          f"(p90 {np.percentile(first, 90):.2f}) | first S->O done per tile {np.median(body / tiles):.3f} us | epilogue {np.median(epi):.2f}")
    a, b = np.polyfit(tiles, total, 1)
    print(f"   CTA time ~= {a:.3f} us * tiles + {b:.2f} us (least squares over CTAs)")
    gaps = []
    for sm in np.unique(r[:, 0]):
        rows = r[r[:, 0] == sm]
        rows = rows[np.argsort(rows[:, 1])]
        gaps += list(us(rows[1:, 1] - rows[:-1, 6]))
    gaps = np.array(gaps)
    print(f"   same-SM gap (stores of CTA i -> entry of CTA i+1): median {np.median(gaps):.2f} us, p90 {np.percentile(gaps, 90):.2f}, "
          f"CTAs per SM {len(r) / len(np.unique(r[:, 0])):.1f}")
