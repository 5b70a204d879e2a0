# 1 -> 8 GPU scaling on ONE box: sequence-sharded C2 and C5, KV-head-sharded C3, reference arm.
mkdir -p gpurun_out
nvidia-smi -L | head -8
for N in 1 2 4 8; do
  if [ $N = 1 ]; then TR="python"; else TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N))"; fi
  timeout 200 $TR bench.py --gpus $N --steps 1000 --warmup 20 --no-cpu-baseline > gpurun_out/scale_c2_n$N.json 2> gpurun_out/scale_c2_n$N.err; echo "c2 n$N rc=$?"
  timeout 200 $TR bench.py --gpus $N --steps 500 --warmup 20 --no-cpu-baseline --workload c5 > gpurun_out/scale_c5_n$N.json 2> gpurun_out/scale_c5_n$N.err; echo "c5 n$N rc=$?"
  if [ $N != 1 ]; then
    timeout 200 $TR bench.py --gpus $N --steps 1000 --warmup 20 --no-cpu-baseline --workload c3 --shard head > gpurun_out/scale_c3head_n$N.json 2> gpurun_out/scale_c3head_n$N.err; echo "c3head n$N rc=$?"
  else
    timeout 200 $TR bench.py --gpus 1 --steps 1000 --warmup 20 --no-cpu-baseline --workload c3 > gpurun_out/scale_c3head_n1.json 2> gpurun_out/scale_c3head_n1.err
  fi
done
python - <<PY
import json
for w in ("c2","c5","c3head"):
    base=None
    for n in (1,2,4,8):
        try:
            d=json.load(open(f"gpurun_out/scale_{w}_n{n}.json"))
            base = base or d["value"]
            print(w, n, round(d["value"]), "tok/s", round(d["ms_per_step"]*1e3,1), "us", d.get("pct_of_8tbs_per_gpu"), "x%.2f"%(d["value"]/base), d["e2e"]["value"] if d.get("e2e") else None)
        except Exception as e: print(w, n, "ERR", e)
PY
