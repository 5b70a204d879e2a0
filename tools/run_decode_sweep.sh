#!/bin/bash
# gpurun helper: decode parity tests, then the tuning sweep + per-CTA timeline of the C2 decode kernel
tag=${1:-sweep}; combos=${2:-16:3:0:40}; wl=${3:-c2}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_decode_strict.py tests/test_gpu_parity.py tests/test_robustness.py -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_pytest.log
tail -4 gpurun_out/${tag}_pytest.log
timeout 600 python tools/exp_balance.py $wl --combos $combos --out gpurun_out/${tag}_timeline.json > gpurun_out/${tag}_sweep.log 2>&1
grep -E "units_per_sm|consumers_end|finisher_end|tickets_ex|first_block|cons_wait" gpurun_out/${tag}_sweep.log
