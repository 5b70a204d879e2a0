#!/bin/bash
# gpurun helper: parity of the generic decode paths + the fp32 C2 roofline point
tag=${1:-gen}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_decode_strict.py tests/test_gpu_parity.py tests/test_engine_adapter.py tests/test_engine_golden.py tests/test_step_tables.py -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_pytest.log
tail -6 gpurun_out/${tag}_pytest.log
timeout 300 python bench.py --steps 200 --warmup 5 --no-cpu-baseline > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err
python - <<PY
import json
d=json.load(open("gpurun_out/${tag}_bench.json"))
print("c2", d["ms_per_step"], d["roofline"]["frac"])
print("generic_f32", json.dumps(d.get("generic_f32")))
PY
