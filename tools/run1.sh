set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
for w in c2 c3 c5; do timeout 300 python tools/exp_balance.py $w > gpurun_out/exp_balance_$w.log 2>&1; cat gpurun_out/exp_balance_$w.log; done
