#!/bin/bash
# gpurun helper: decode parity tests, C2 / C3 timelines and the per-rank head-sharded shapes - everything under a timeout
tag=${1:-da}; combos=${2:-16:3:0:32}
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_decode_strict.py tests/test_gpu_parity.py tests/test_robustness.py tests/test_peer_gather.py -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_pytest.log
tail -3 gpurun_out/${tag}_pytest.log
for wl in c2 c3 c5; do
  timeout 120 python tools/exp_balance.py $wl --combos $combos --out gpurun_out/${tag}_timeline_$wl.json > gpurun_out/${tag}_sweep_$wl.log 2>&1; echo "$wl rc=$?"
  grep -E "units_per_sm|consumers_end|finisher_end" gpurun_out/${tag}_sweep_$wl.log
done
rm -f gpurun_out/hs.log
bash tools/run_hs.sh "8 4 2" 16:3:0:32
