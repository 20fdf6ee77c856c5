#!/bin/bash
O=gpurun_out
mkdir -p $O
for t in 1 4 8 16; do ./tests/cpp/_bin/bench_search gpu_op_mt 0 1 4096 6 1 $t; done > $O/r2v_search_op_mt.jsonl 2>&1
./tests/cpp/_bin/bench_search ref_op_mt 0 1 4096 3 1 16 >> $O/r2v_search_op_mt.jsonl 2>&1
./tests/cpp/_bin/bench_search gpu 0 1 4096 6 1 >> $O/r2v_search_op_mt.jsonl 2>&1
./tests/cpp/_bin/bench_search ref 0 1 4096 6 1 >> $O/r2v_search_op_mt.jsonl 2>&1
cut -c1-260 $O/r2v_search_op_mt.jsonl
timeout 600 python bench.py --steps 5 --warmup 3 > $O/r2v_bench_full.json 2> $O/r2v_bench_full.err; python tools/show_bench.py $O/r2v_bench_full.json | tail -6; tail -2 $O/r2v_bench_full.err
