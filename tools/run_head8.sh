# KV-head-sharded C3 (8 x 32k) strong scaling on one box: fused peer gather (LL / counters, graphs) vs NCCL all-gather
mkdir -p gpurun_out
timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29530 tools/peer_gather_check.py 2>&1 | grep "peer_gather_check\|Error\|error\|diff" | tail -12
timeout 200 python bench.py --gpus 1 --steps 1000 --warmup 20 --no-cpu-baseline --workload c3 > gpurun_out/head_n1.json 2> gpurun_out/head_n1.err
for N in 4 8; do
  TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29540 + N))"
  for G in peer peer-counter nccl; do
    timeout 200 $TR bench.py --gpus $N --steps 1000 --warmup 20 --no-cpu-baseline --workload c3 --shard head --gather $G > gpurun_out/head_${G}_n$N.json 2> gpurun_out/head_${G}_n$N.err; echo "$G n$N rc=$?"
  done
done
python - <<PY
import json
base=None
for f in ["head_n1"]+[f"head_{g}_n{n}" for n in (4,8) for g in ("peer","peer-counter","nccl")]:
    try:
        d=json.loads([l for l in open(f"gpurun_out/{f}.json") if l.startswith("{")][0])
        base = base or d["value"]
        print(f, round(d["value"]), "tok/s", round(d["ms_per_step"]*1e3,1), "us/step", "x%.2f" % (d["value"]/base), "kernel", round(d["roofline"]["kernel_ms"]*1e3,1), "e2e", round(d["e2e"]["value"]), round(d["e2e"]["ms_per_step"]*1e3,1),"us")
    except Exception as e: print(f, "ERR", e)
PY
