# Re-verification run: GPU tests, C2/C3/C5 bench lines, prefill bench, launch list and ncu captures.
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "c2 rc=$?"
timeout 300 python bench.py --workload c3 --no-cpu-baseline --steps 1000 > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err
timeout 300 python bench.py --workload c5 --no-cpu-baseline --steps 500 > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
timeout 300 python tools/bench_prefill.py > gpurun_out/bench_prefill.jsonl 2> gpurun_out/bench_prefill.err
cat gpurun_out/bench_prefill.jsonl | cut -c1-400
PF="python tools/bench_prefill.py --shape full8k --iters 3 --warmup 2"
$PF > gpurun_out/pf_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:pa_prefill -s 2 -c 1 -o gpurun_out/prof_prefill_tc2 $PF > gpurun_out/ncu_prefill.log 2>&1
tail -3 gpurun_out/ncu_prefill.log
head -c 2500 gpurun_out/bench_c2.json; echo; head -c 500 gpurun_out/bench_c3.json; echo; head -c 500 gpurun_out/bench_c5.json; echo; head -c 600 gpurun_out/bench_ref.json; tail -5 gpurun_out/bench_c2.err
