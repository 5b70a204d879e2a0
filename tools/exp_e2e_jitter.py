"""Experiment: how steady is the end-to-end decode step (bench.py's e2e leg), and does the NVML clock sampler
disturb it?  Per-step host wall clock over N steps, with the sampler off / at 2 ms / at 20 ms: mean, median,
p99, max and the number of steps above twice the median (one 14 ms hiccup in a 300-step region moves the
mean from 120 to 168 us)."""
import os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from vllm_from_scratch_b200.graph import DecodeStep

dev = torch.device("cuda", 0)
wl = bench.WORKLOADS["c2"]
b, ctx, hq, hkv, d, bs, layers, _ = wl
layers = 4
pool, bts_h, lens_h, q_h, k_h, v_h, slots_h = bench.make_workload(torch, dev, wl, layers, seed=1)
steps = []
for i in range(layers):
    st = DecodeStep(pool.layer(i), b, hq, bts_h.shape[1], write_kv=True)
    st.host["q"].copy_(q_h); st.host["k"].copy_(k_h); st.host["v"].copy_(v_h); st.host["slots"].copy_(slots_h)
    st.host["block_tables"].copy_(bts_h); st.host["seq_lens"].copy_(lens_h)
    steps.append(st.capture())


def run(name, n):
    for i in range(20):
        steps[i % layers].replay(); steps[i % layers].wait()
    torch.cuda.synchronize()
    ts = []
    t = time.perf_counter()
    for i in range(n):
        s = steps[i % layers]
        s.replay()
        s.wait()
        t2 = time.perf_counter()
        ts.append((t2 - t) * 1e6)
        t = t2
    mean = sum(ts) / n
    srt = sorted(ts)
    med = srt[n // 2]
    slow = [(i, round(x)) for i, x in enumerate(ts) if x > 2 * med]
    print(f"{name:28s} n={n:5d} mean {mean:7.1f} median {med:7.1f} p99 {srt[int(n * 0.99)]:7.1f} max {srt[-1]:9.1f} us; "
          f"steps > 2x median: {len(slow)} {slow[:6]}", flush=True)


for rep in range(2):
    run("sampler off", 300)
    run("sampler off", 2000)
    for period in (0.002, 0.02):
        sm = bench.ClockSampler(0, period_s=period)
        sm.start()
        run(f"sampler every {period * 1e3:.0f} ms", 300)
        run(f"sampler every {period * 1e3:.0f} ms", 2000)
        sm.stop_flag.set()
        sm.join()
