"""Bring-up aid for the tcgen05 prefill kernel: runs a few small shapes and prints WHERE the
output differs from the oracle (which rows / heads / columns), which tells a descriptor or
swizzle mistake from a pipeline one.  Not a test; `python tools/prefill_debug.py` on a B200."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import paged_attention_oracle as O                      # noqa: E402
from vllm_from_scratch_b200 import paged_attention as P             # noqa: E402


def run(total, hq, hkv, d, start=0, seed=0, ident=False):
    g = torch.Generator().manual_seed(seed)
    bs, nb = 16, (total + 15) // 16 + 4
    q = torch.randn(total, hq, d, generator=g).to(torch.bfloat16)
    k = torch.randn(total, hkv, d, generator=g).to(torch.bfloat16)
    v = torch.randn(total, hkv, d, generator=g).to(torch.bfloat16)
    if ident:       # q = 0 -> uniform softmax: isolates the PV half
        q.zero_()
    bt = torch.randperm(nb, generator=g)[:(total + bs - 1) // bs].tolist()
    kv = torch.zeros(2, nb, hkv, bs, d, dtype=torch.bfloat16)
    O.write_kv_hbsd(kv, bt, k, v, bs)
    bt_dev = torch.tensor(bt, dtype=torch.int32, device="cuda")
    try:
        got = P.paged_attention_prefill_chunk(q[start:].cuda(), kv.cuda(), bt_dev, start)
        torch.cuda.synchronize()
    except Exception as e:                                            # noqa: BLE001
        print(f"total {total} hq {hq} hkv {hkv} d {d} start {start}: FAILED {e}")
        return False
    ref = O.chunked_prefill_hbsd(q[start:], kv, bt, start, bs, compute_dtype=torch.float64)
    err = (got.double().cpu() - ref).abs()
    bad = err > 2e-2
    print(f"total {total} hq {hq} hkv {hkv} d {d} start {start} ident {ident}: max err {err.max().item():.4g} "
          f"nan {int(torch.isnan(got).sum())} bad {int(bad.sum())}/{bad.numel()}")
    if bad.any() or torch.isnan(got).any():
        bad = bad | torch.isnan(got).cpu()
        print("  bad tokens :", bad.any(2).any(1).nonzero().flatten().tolist()[:40])
        print("  bad heads  :", bad.any(2).any(0).nonzero().flatten().tolist())
        print("  bad columns:", bad.any(1).any(0).nonzero().flatten().tolist()[:128])
        t0 = int(bad.any(2).any(1).nonzero().flatten()[0])
        print("  got[t0,0,:8]", got[t0, 0, :8].float().cpu().tolist())
        print("  ref[t0,0,:8]", ref[t0, 0, :8].tolist())
        return False
    return True


if __name__ == "__main__":
    ok = True
    for ident in (True, False):
        ok &= run(16, 4, 4, 64, ident=ident)
        ok &= run(16, 4, 4, 128, ident=ident)
        ok &= run(128, 4, 4, 128, ident=ident)
    ok &= run(128, 32, 8, 128)
    ok &= run(300, 32, 8, 128)
    ok &= run(300, 32, 8, 128, start=100)
    ok &= run(1000, 8, 2, 64)
    print("ALL OK" if ok else "SOME FAILED")
