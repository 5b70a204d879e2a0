set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv -lms 200 > gpurun_out/clocks.csv &
SMI=$!
timeout 600 python bench.py > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err
timeout 300 python bench.py --workload c3 --no-cpu-baseline --steps 1000 > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err
timeout 300 python bench.py --workload c5 --no-cpu-baseline --steps 500 > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err
kill $SMI
SHORT="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --ramp-ms 0 --e2e-steps 5"
$SHORT > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $SHORT > gpurun_out/ncu_launches.log 2>&1
$SHORT > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:pa_decode_ws -s 10 -c 3 -o gpurun_out/prof_c2 $SHORT > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
head -c 3000 gpurun_out/bench_c2.json; echo; head -c 600 gpurun_out/bench_c3.json; echo; head -c 600 gpurun_out/bench_c5.json; cat gpurun_out/bench_c2.err | tail -5
