mkdir -p gpurun_out
N=${N:-2}
export BENCH_DEBUG=1 BENCH_DEBUG_AFTER=100
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 150 $TR bench.py --gpus $N --steps 1000 --warmup 20 --no-cpu-baseline > gpurun_out/A.json 2> gpurun_out/A.err; echo "A rc=$? $(wc -c < gpurun_out/A.json)"
timeout 150 $TR bench.py --gpus $N --steps 200 --warmup 5 > gpurun_out/B.json 2> gpurun_out/B.err; echo "B rc=$? $(wc -c < gpurun_out/B.json)"
timeout 150 $TR bench.py --gpus $N --steps 200 --warmup 5 --no-cpu-baseline > gpurun_out/C.json 2> gpurun_out/C.err; echo "C rc=$? $(wc -c < gpurun_out/C.json)"
timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 --tee 3 --log-dir gpurun_out/trlogs bench.py --gpus $N --steps 200 --warmup 5 > gpurun_out/D.json 2> gpurun_out/D.err; echo "D rc=$? $(wc -c < gpurun_out/D.json)"
for f in A B C D; do echo "== $f"; grep -v "^\*\*\*\|OMP_NUM" gpurun_out/$f.err | cut -c1-200 | tail -12; done
