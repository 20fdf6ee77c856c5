#!/bin/bash
# full GPU suite + full bench on the build with k_warp2
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/r2s_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/r2s_pytest_gpu.log; tail -4 $O/r2s_pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 > $O/r2s_bench_full.json 2> $O/r2s_bench_full.err; python tools/show_bench.py $O/r2s_bench_full.json; tail -2 $O/r2s_bench_full.err
timeout 300 python tests/fuzz_parity.py 600 31 > $O/r2s_fuzz_600_seed31.log 2>&1; tail -4 $O/r2s_fuzz_600_seed31.log
