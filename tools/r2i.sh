#!/bin/bash
# round 2: eight GPUs — rank-per-GPU bench (all workloads), the vector-store variant, the single-process context, the batched search over all GPUs
O=gpurun_out
mkdir -p $O
nvidia-smi -L > $O/r2i_gpus.txt; nproc > $O/r2i_nproc.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29521 bench.py --gpus 8 --steps 10 --warmup 3 > $O/r2i_bench_8gpu.json 2> $O/r2i_bench_8gpu.err
tail -c 300 $O/r2i_bench_8gpu.err
python tools/show_bench.py $O/r2i_bench_8gpu.json
for rb in 64 128; do
  timeout 300 $TR --master-port 2953$((rb/64)) bench.py --gpus 8 --steps 10 --warmup 3 --no-extra --no-cpu --rb $rb > $O/r2i_bench_8gpu_rb$rb.json 2> $O/r2i_bench_8gpu_rb$rb.err
  python tools/show_bench.py $O/r2i_bench_8gpu_rb$rb.json
done
T=$(nproc)
tests/cpp/_bin/bench_search gpu_tb_batched -1 0 4096 150 0 $T $((3*4096*8/T)) > $O/r2i_search_8gpu.jsonl 2> $O/r2i_search_8gpu.err
tests/cpp/_bin/bench_search ref_tb_mt 0 0 0 150 0 $T 64 >> $O/r2i_search_8gpu.jsonl 2>> $O/r2i_search_8gpu.err
tests/cpp/_bin/bench_search ref_tb 0 0 0 20000 2000 >> $O/r2i_search_8gpu.jsonl 2>> $O/r2i_search_8gpu.err
cut -c1-300 $O/r2i_search_8gpu.jsonl
