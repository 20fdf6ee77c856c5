#!/bin/bash
# round 2: two GPUs — the multi-GPU context with a real NCCL all-reduce, and the rank-per-GPU bench
O=gpurun_out
mkdir -p $O
nvidia-smi -L > $O/r2g_gpus.txt
NCCL_DEBUG=WARN timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_dropin_cpp.py -m gpu -x -q -k "multi or words_mode_mold" > $O/r2g_pytest.log 2>&1; echo "pytest rc=$?" >> $O/r2g_pytest.log; tail -6 $O/r2g_pytest.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > $O/r2g_bench_2gpu.json 2> $O/r2g_bench_2gpu.err
tail -c 300 $O/r2g_bench_2gpu.err
python tools/show_bench.py $O/r2g_bench_2gpu.json
