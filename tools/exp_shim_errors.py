"""Experiment: which error paths of the torch shim survive once a CUDA context is up (each case in its own process)."""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CASES = {
 "raise3_no_cuda": "ext._raise(3, 7)",
 "raise3_after_cuda_init": "torch.zeros(1, device='cuda'); ext._raise(3, 7)",
 "raise3_after_decode": "P.paged_attention_decode_batched(q, kv, bts, sl); torch.cuda.synchronize(); ext._raise(3, 7)",
 "raise0_after_decode": "P.paged_attention_decode_batched(q, kv, bts, sl); torch.cuda.synchronize(); ext._raise(0, 7)",
 "prefill_chunk_short": "P.paged_attention_prefill_chunk(qq, kvb, btb[0, :3], 30)",
 "decode_bad_dtype": "P.paged_attention_decode_batched(q, kv, bts.long(), sl)",
}
PRE = '''
import faulthandler, sys, torch
faulthandler.enable()
sys.path.insert(0, %r)
from vllm_from_scratch_b200 import _lib, paged_attention as P
ext = _lib.load_ext()
if %r:
    q = torch.randn(2, 4, 32, device="cuda"); kv = torch.randn(2, 8, 4, 16, 32, device="cuda")
    bts = torch.zeros(2, 2, dtype=torch.int32, device="cuda"); sl = torch.tensor([5, 20], dtype=torch.int32, device="cuda")
    qq = torch.randn(20, 32, 128, device="cuda").to(torch.bfloat16); kvb = torch.randn(2, 40, 8, 16, 128, device="cuda").to(torch.bfloat16)
    btb = torch.arange(12, dtype=torch.int32, device="cuda").view(3, 4)
try:
    %s
    print("no exception")
except Exception as e:
    print(type(e).__name__, str(e)[:80])
'''
for name, body in CASES.items():
    r = subprocess.run([sys.executable, "-c", PRE % (ROOT, "no_cuda" not in name, body)], capture_output=True, text=True, timeout=120)
    print(f"{name:26s} rc={r.returncode} {r.stdout.strip()[:120]} {'SEGV' if r.returncode < 0 else ''}", flush=True)
