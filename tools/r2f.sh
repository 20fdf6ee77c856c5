#!/bin/bash
# round 2, second GPU call: parity suite incl. mode W / batching adapter / callers, search benchmarks, full bench line
mkdir -p gpurun_out
nproc > gpurun_out/r2f_nproc.txt
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r2f_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2f_pytest.log
tail -15 gpurun_out/r2f_pytest.log
B=tests/cpp/_bin/bench_search
T=$(nproc)
{
  $B ref_tb 0 0 0 20000 2000
  $B gpu_tb 0 0 0 5000 500
  $B ref_tb_mt 0 0 0 150 0 $T 64
  $B gpu_tb_batched 0 0 4096 150 0 $T 768
  $B gpu_tb_batched 0 0 4096 150 0 $T 256
  $B gpu_tb_batched 0 0 1024 150 0 $T 256
  $B gpu_tb_batched 0 0 4096 150 0 $((T/2)) 1536
  $B gpu_tb_batched 0 0 4096 600 0 $T 768
} > gpurun_out/r2f_search.jsonl 2> gpurun_out/r2f_search.err
cat gpurun_out/r2f_search.jsonl | cut -c1-330
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r2f_bench_full.json 2> gpurun_out/r2f_bench_full.err
tail -c 400 gpurun_out/r2f_bench_full.err
python tools/show_bench.py gpurun_out/r2f_bench_full.json 2>&1 | tail -20
