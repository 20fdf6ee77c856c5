#!/bin/bash
# round 2: k_warp2 — parity of the warp path, then the M2-DR bench line
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "warp2 or extreme or general_path or lane_boundary or words_mode or private" > $O/r2l_pytest_warp.log 2>&1; echo "pytest rc=$?" >> $O/r2l_pytest_warp.log; tail -5 $O/r2l_pytest_warp.log
timeout 300 python bench.py --workload tree_d20_9x9x3_dr --steps 5 --warmup 3 --no-cpu --no-extra > $O/r2l_bench_m2dr.json 2> $O/r2l_bench_m2dr.err; python tools/show_bench.py $O/r2l_bench_m2dr.json; tail -3 $O/r2l_bench_m2dr.err
