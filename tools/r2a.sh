#!/bin/bash
# round 2, first GPU call: parity suite on the refactored library, then a bench line and an rb sweep
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r2a_gpus.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2a_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2a_pytest.log
tail -5 gpurun_out/r2a_pytest.log
for rb in 32 64 128; do
  timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-extra --rb $rb > gpurun_out/r2a_bench_A_rb$rb.json 2> gpurun_out/r2a_bench_A_rb$rb.err
done
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r2a_bench_full.json 2> gpurun_out/r2a_bench_full.err
tail -c 600 gpurun_out/r2a_bench_full.err
python tools/show_bench.py gpurun_out/r2a_bench_A_rb32.json gpurun_out/r2a_bench_A_rb64.json gpurun_out/r2a_bench_A_rb128.json gpurun_out/r2a_bench_full.json 2>&1 | tail -20
