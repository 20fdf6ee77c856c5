#!/bin/bash
# round 2, final build on eight GPUs: multi-GPU parity tests, rank-per-GPU bench at 8 (all workloads), 4 and 2 (headline only)
O=gpurun_out
mkdir -p $O
nvidia-smi -L | wc -l > $O/r2y_gpus.txt
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "multi" > $O/r2y_pytest_multi.log 2>&1; echo "pytest rc=$?" >> $O/r2y_pytest_multi.log; tail -2 $O/r2y_pytest_multi.log
timeout 300 ./tests/cpp/_bin/test_dropin --multi 8 > $O/r2y_dropin_multi8.log 2>&1; echo "dropin rc=$?" >> $O/r2y_dropin_multi8.log; tail -3 $O/r2y_dropin_multi8.log
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 900 $TR --nproc-per-node 8 --master-port 29521 bench.py --gpus 8 --steps 10 --warmup 3 > $O/r2y_bench_8gpu.json 2> $O/r2y_bench_8gpu.err
tail -c 300 $O/r2y_bench_8gpu.err
python tools/show_bench.py $O/r2y_bench_8gpu.json
for n in 4 2; do
  timeout 300 $TR --nproc-per-node $n --master-port 2953$n bench.py --gpus $n --steps 10 --warmup 3 --no-extra --no-cpu > $O/r2y_bench_${n}gpu.json 2> $O/r2y_bench_${n}gpu.err
  python tools/show_bench.py $O/r2y_bench_${n}gpu.json | head -2
done
