#!/bin/bash
# gpurun helper: the engine-loop tests + the C4 engine bench
tag=${1:-engine}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_step_tables.py tests/test_engine_adapter.py tests/test_engine_golden.py tests/test_kv_manager.py tests/test_graph_step.py -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_pytest.log
tail -30 gpurun_out/${tag}_pytest.log
timeout 300 python tools/bench_engine.py > gpurun_out/${tag}_bench_engine.jsonl 2> gpurun_out/${tag}_bench_engine.err; cat gpurun_out/${tag}_bench_engine.jsonl; tail -5 gpurun_out/${tag}_bench_engine.err
