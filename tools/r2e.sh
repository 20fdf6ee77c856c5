#!/bin/bash
# round 2: paired generator halves + lean consumer: parity subset, sweeps, ncu
O=gpurun_out
mkdir -p $O
timeout 1200 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "shared_root or geometry or full_size or mold or known_answers or packed or multi or continuation or jump_ahead or zero_copy" > $O/r2e_pytest.log 2>&1; echo "pytest rc=$?" >> $O/r2e_pytest.log; tail -8 $O/r2e_pytest.log
python tools/sweep_stream.py A B M > $O/r2e_sweep.log 2>&1; cat $O/r2e_sweep.log
for rb in 32 64 128; do PCU_RB=$rb python tools/one_call.py A 8 ; done > $O/r2e_rb.log 2>&1; cat $O/r2e_rb.log
for s in 8; do
  ncu --set full --import-source on --clock-control none -k regex:"k_stream2" -c 1 -o $O/r2e_stream_s$s -f python tools/one_call.py A $s > $O/r2e_ncu_$s.log 2>&1
  ncu -i $O/r2e_stream_s$s.ncu-rep --page raw --csv > $O/r2e_stream_s$s.csv 2>/dev/null
done
python tools/ncu_summary.py $O/r2e_stream_s8.csv > $O/r2e_ncu_summary.csv
cat $O/r2e_ncu_summary.csv | cut -c1-1200
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-extra > $O/r2e_bench_A.json 2> $O/r2e_bench_A.err; python tools/show_bench.py $O/r2e_bench_A.json
