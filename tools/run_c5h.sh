# C5's 32k shape (64 x 32768, 8 GiB of KV per layer): 1 GPU vs KV-head-sharded over 8 GPUs
mkdir -p gpurun_out
timeout 300 python bench.py --gpus 1 --steps 100 --warmup 5 --no-cpu-baseline --workload c5h > gpurun_out/c5h_n1.json 2> gpurun_out/c5h_n1.err; echo "n1 rc=$?"
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29561"
for G in peer nccl; do
  timeout 300 $TR bench.py --gpus 8 --steps 300 --warmup 10 --no-cpu-baseline --workload c5h --shard head --gather $G > gpurun_out/c5h_${G}_n8.json 2> gpurun_out/c5h_${G}_n8.err; echo "$G rc=$?"
done
tail -3 gpurun_out/c5h_n1.err | cut -c1-200
python - <<PY
import json
base=None
for f in ("c5h_n1","c5h_peer_n8","c5h_nccl_n8"):
    try:
        d=json.loads([l for l in open(f"gpurun_out/{f}.json") if l.startswith("{")][0])
        base = base or d["value"]
        print(f, round(d["value"]), "tok/s", round(d["ms_per_step"]*1e3,1), "us/step", "x%.2f" % (d["value"]/base), "kernel", round(d["roofline"]["kernel_ms"]*1e3,1), "pct8", round(d["pct_of_8tbs_per_gpu"],1), "e2e", round(d["e2e"]["value"]))
    except Exception as e: print(f, "ERR", e)
PY
