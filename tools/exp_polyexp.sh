# Experiment: share of the softmax exponentials sent through the FMA-pipe polynomial (kPolyExpMask)
set -e
F=vllm_from_scratch_b200/csrc/pa_prefill_tc2.cu
cp $F /tmp/tc2.bak
for M in 0x00u 0x22u 0x4Au 0xAAu 0xDBu; do
  sed -i "s/constexpr uint32_t kPolyExpMask = 0x[0-9A-Fa-f]*u;/constexpr uint32_t kPolyExpMask = $M;/" $F
  python -m vllm_from_scratch_b200.build > /dev/null
  echo "== mask $M"
  timeout 200 python tools/bench_prefill.py 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print('  ', d['shape'], round(d['ms'], 3), 'ms', round(d['tflops'], 1), 'TF/s')"
done
cp /tmp/tc2.bak $F
python -m vllm_from_scratch_b200.build > /dev/null
timeout 300 python -m pytest tests/test_prefill_tc.py tests/test_engine_adapter.py -m gpu -x -q 2>&1 | tail -3
