set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -25 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --no-cpu-baseline --steps 1000 > gpurun_out/bench_c2_graph.json 2> gpurun_out/bench_c2_graph.err; tail -3 gpurun_out/bench_c2_graph.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_c2_graph.json'))
print({k:d[k] for k in ('value','ms_per_step','pct_of_8tbs_per_gpu')}, d['e2e'], d['roofline']['frac'])
PY
