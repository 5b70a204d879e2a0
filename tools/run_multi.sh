mkdir -p gpurun_out
N=${N:-2}
export BENCH_DEBUG=1 BENCH_DEBUG_AFTER=100
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 150 $TR bench.py --gpus $N --steps 1000 --warmup 20 > gpurun_out/bench_seq_n$N.json 2> gpurun_out/bench_seq_n$N.err; echo "seq rc=$?"
timeout 150 $TR bench.py --gpus $N --steps 1000 --warmup 20 --shard head > gpurun_out/bench_head_n$N.json 2> gpurun_out/bench_head_n$N.err; echo "head rc=$?"
timeout 150 $TR bench.py --gpus $N --steps 500 --warmup 20 --workload c5 > gpurun_out/bench_c5_n$N.json 2> gpurun_out/bench_c5_n$N.err; echo "c5 rc=$?"
timeout 150 $TR bench.py --gpus $N --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_n$N.json 2> gpurun_out/bench_ref_n$N.err; echo "ref rc=$?"
for f in seq head c5; do echo "== $f"; grep -v "^\*\*\*\|OMP_NUM" gpurun_out/bench_${f}_n$N.err | cut -c1-200 | tail -40; done
python - <<PY
import json
for f in ("seq","head","c5","ref"):
    try:
        d=json.load(open(f"gpurun_out/bench_{f}_n$N.json"))
        print(f, {k:d.get(k) for k in ('value','ms_per_step','n_gpus','scaling','pct_of_8tbs_per_gpu')}, d.get('e2e',{}).get('value'))
    except Exception as e: print(f, "ERR", e)
PY
