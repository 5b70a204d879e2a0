mkdir -p gpurun_out
N=${N:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521"
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 120 $TR tools/peer_gather_check.py 2>&1 | grep "peer_gather_check\|Error\|error\|diff" | tail -12
for G in peer peer-counter nccl; do
  timeout 200 $TR bench.py --gpus $N --steps 1000 --warmup 20 --no-cpu-baseline --workload c3 --shard head --gather $G > gpurun_out/head_${G}_n$N.json 2> gpurun_out/head_${G}_n$N.err; echo "$G rc=$?"
done
timeout 200 python bench.py --steps 1000 --warmup 20 --no-cpu-baseline > gpurun_out/c2_after_ll.json 2> gpurun_out/c2_after_ll.err
grep -v "^\*\*\*\|OMP_NUM\|^$" gpurun_out/head_peer_n$N.err | tail -6
python - <<PY
import json
for f in ("head_peer_n$N","head_peer-counter_n$N","head_nccl_n$N","c2_after_ll"):
    try:
        d=json.loads([l for l in open(f"gpurun_out/{f}.json") if l.startswith("{")][0])
        print(f, round(d["value"]), "tok/s", round(d["ms_per_step"]*1e3,1), "us/step", "kernel", round(d["roofline"]["kernel_ms"]*1e3,1), "e2e", round(d["e2e"]["value"]), round(d["e2e"]["ms_per_step"]*1e3,1),"us")
    except Exception as e: print(f, "ERR", e)
PY
