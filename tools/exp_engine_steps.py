"""Experiment: per-step wall time of the C4 engine loop (tools/bench_engine.py workload) with decode graphs."""
import os, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "tools"))
from engine_harness import TinyEngine
from vllm_from_scratch_b200 import engine_adapter, kv_manager, _lib, paged_attention as P
import bench_engine

pdl = int(sys.argv[1]) if len(sys.argv) > 1 else 1
dtype = torch.bfloat16 if (len(sys.argv) < 3 or sys.argv[2] == "bf16") else torch.float32
_lib.load().pa_decode_set_pdl(pdl)
hidden = 1024 if dtype == torch.bfloat16 else 256
eng = TinyEngine(vocab=1000, hidden=hidden, layers=4, heads=8, inter=4 * hidden, num_blocks=40, block_size=16,
                 max_seqs=16, token_budget=64, device="cuda", dtype=dtype, hashed_blocks=True)
engine_adapter.install(eng, ops=P, cuda_graphs=True)
kv_manager.install(eng, num_host_blocks=256, ops=P)
ps = bench_engine.prompts()
for rnd in range(2):
    reqs = [eng.add_request(p, max_new_tokens=20) for p in ps]
    times = []
    g = eng._pa_b200["graphs"]
    while eng.scheduler.has_unfinished:
        r0, c0 = g.replays, g.captures
        torch.cuda.synchronize(); t0 = time.perf_counter()
        s = eng.step()
        torch.cuda.synchronize(); dt = time.perf_counter() - t0
        if not s and not eng.scheduler.paused:
            break
        md = eng._pa_b200["meta"]
        times.append((dt * 1e3, g.replays - r0, g.captures - c0, len(s), sum(len(x.tokens) for x in s)))
    print(f"round {rnd} pdl={pdl}: total {sum(t[0] for t in times):.1f} ms")
    for t in times:
        print("   %.2f ms replay=%d capture=%d reqs=%d tokens=%d" % t)
