#!/bin/bash
# gpurun --gpus 8 helper: the driver-style bench at N = 8 (sequence-sharded C2 headline + head_sharded extras) and the
# multi-rank parity test; everything under a timeout
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29508"
timeout 300 $TR bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/n8_bench_s20.json 2> gpurun_out/n8_bench_s20.err; echo "bench s20 rc=$?"
timeout 300 $TR bench.py --gpus 8 --steps 1000 --warmup 20 --no-extras --no-cpu-baseline > gpurun_out/n8_bench_s1000.json 2> gpurun_out/n8_bench_s1000.err; echo "bench s1000 rc=$?"
timeout 400 python -m pytest tests/test_multi_gpu.py -m gpu -x -q 2>&1 | tail -3
python - <<PY
import json
for f in ("n8_bench_s20", "n8_bench_s1000"):
    try:
        d = json.load(open(f"gpurun_out/{f}.json"))
        print(f, round(d["value"]), "tok/s", round(d["ms_per_step"] * 1e3, 2), "us", "e2e", round(d["e2e"]["value"]) if d.get("e2e") else None, d["clocks"]["sm_mhz"])
        hs = d.get("head_sharded") or {}
        for k, v in hs.items():
            print("  ", k, "one_gpu_us", round(v["one_gpu_us_per_step"], 1), {a: (round(x["us_per_step"], 1), round(x["speedup_vs_one_gpu"], 2), x["parity_checked"]) for a, x in v["arms"].items()})
    except Exception as e:
        print(f, "ERR", e)
PY
