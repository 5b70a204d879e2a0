#!/bin/bash
# gpurun helper: per-rank head-sharded shapes, every configuration under its own timeout (a hang must not eat the box)
mkdir -p gpurun_out
for w in ${1:-2 8 4 1}; do
  echo "== WORLDS=$w" >> gpurun_out/hs.log
  WORLDS=$w COMBOS=${2:-16:3:0:32,64:3:0:32} timeout 90 python tools/exp_headshard.py >> gpurun_out/hs.log 2>&1; echo "rc=$?" >> gpurun_out/hs.log
done
grep -E "==|rc=|units_per|fused|per-CTA" gpurun_out/hs.log
