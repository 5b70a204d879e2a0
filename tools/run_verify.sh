#!/bin/bash
# One gpurun call: GPU parity suite, smoke, the default bench line and the driver-style short bench.
# Usage: gpurun --timeout 1500 -- 'bash tools/run_verify.sh [tag]'
tag=${1:-verify}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/${tag}_gpus.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_pytest.log
tail -5 gpurun_out/${tag}_pytest.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/${tag}_smoke.log 2>&1; tail -2 gpurun_out/${tag}_smoke.log
timeout 600 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; tail -c 3000 gpurun_out/${tag}_bench.json
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/${tag}_bench_s20.json 2> gpurun_out/${tag}_bench_s20.err
