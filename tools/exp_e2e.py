"""Experiment: where the end-to-end decode step (graph.DecodeStep, bench.py's e2e leg) spends its time over the kernel."""
import os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from vllm_from_scratch_b200 import paged_attention as P
from vllm_from_scratch_b200.graph import DecodeStep

dev = torch.device("cuda", 0)
wl = bench.WORKLOADS["c2"]
b, ctx, hq, hkv, d, bs, layers, _ = wl
layers = 4
pool, bts_h, lens_h, q_h, k_h, v_h, slots_h = bench.make_workload(torch, dev, wl, layers, seed=1)
caches = [pool.layer(i) for i in range(layers)]


def run(name, steps, n=300):
    for i in range(20):
        steps[i % layers].replay(); steps[i % layers].wait()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(n):
        s = steps[i % layers]
        s.replay()
        s.wait()
    dt = (time.perf_counter() - t0) / n * 1e6
    print(f"{name:58s} {dt:7.1f} us / step (host wall clock, result waited for every step)", flush=True)


def make(write_kv, zero_copy, out=None):
    res = []
    for c in caches:
        st = DecodeStep(c, b, hq, bts_h.shape[1], write_kv=write_kv, zero_copy=zero_copy, out=out)
        st.host["q"].copy_(q_h)
        if write_kv:
            st.host["k"].copy_(k_h); st.host["v"].copy_(v_h); st.host["slots"].copy_(slots_h)
        st.host["block_tables"].copy_(bts_h); st.host["seq_lens"].copy_(lens_h)
        res.append(st.capture())
    return res


run("staged H2D + slot write + decode, q / out zero-copy (e2e leg)", make(True, True))
run("staged H2D + decode, q / out zero-copy (no slot write)", make(False, True))
run("staged H2D (q too) + slot write + decode + D2H copy of out", make(True, False))
dev_out = torch.empty(b, hq, d, dtype=torch.bfloat16, device=dev)
# decode only, everything resident, host waits every step
q, bts, lens = q_h.to(dev), bts_h.to(dev), lens_h.to(dev)
ws = P.new_decode_workspace(b, hq, hkv, d, dev)
graphs = bench.capture_steps(torch, lambda i: P.paged_attention_decode_batched(q, caches[i], bts, lens, out=dev_out, workspace=ws), layers)
ev = torch.cuda.Event()
for i in range(20):
    graphs[i % layers].replay()
torch.cuda.synchronize()
t0 = time.perf_counter()
for i in range(300):
    graphs[i % layers].replay(); ev.record(); ev.synchronize()
print(f"{'decode kernel only, resident inputs, host waits every step':58s} {(time.perf_counter() - t0) / 300 * 1e6:7.1f} us / step", flush=True)
