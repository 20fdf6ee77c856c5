#!/bin/bash
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "warp2 or extreme or general_path or lane_boundary or words_mode or private" > $O/r2t_pytest_warp.log 2>&1; echo "pytest rc=$?" >> $O/r2t_pytest_warp.log; tail -3 $O/r2t_pytest_warp.log
python tools/one_call_dr.py 20 0 3 0 1 > $O/r2t_one.log 2>&1
for lock in 1 2 4; do python tools/one_call_dr.py 20 1 3 $lock 1 >> $O/r2t_one.log 2>&1; done
for lock in 1 2; do python tools/one_call_dr.py 20 1 3 $lock 0 >> $O/r2t_one.log 2>&1; done
cat $O/r2t_one.log
timeout 300 python bench.py --workload tree_d20_9x9x3_dr --steps 10 --warmup 3 --no-cpu --no-extra > $O/r2t_bench_m2dr.json 2> $O/r2t_bench_m2dr.err; python tools/show_bench.py $O/r2t_bench_m2dr.json | head -2
