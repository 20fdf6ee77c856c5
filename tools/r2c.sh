#!/bin/bash
# round 2, third GPU call: micro-optimised generator (twist mag on the FMA pipe): quick parity, sweep, pipe probes, ncu at 4/8/16 streams per SM
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "shared_root or geometry or full_size or mold_stream or known_answers" > $O/r2c_pytest.log 2>&1; echo "pytest rc=$?" >> $O/r2c_pytest.log; tail -3 $O/r2c_pytest.log
python tools/pipes.py > $O/r2c_pipes.log 2>&1; cat $O/r2c_pipes.log
python tools/sweep_stream.py A > $O/r2c_sweep.log 2>&1; cat $O/r2c_sweep.log
for s in 4 8 16; do
  python tools/one_call.py A $s > $O/r2c_one_$s.log 2>&1; cat $O/r2c_one_$s.log
  ncu --set full --import-source on --clock-control none -k regex:"k_stream2" -c 1 -o $O/r2c_stream_s$s -f python tools/one_call.py A $s > $O/r2c_ncu_$s.log 2>&1
  ncu -i $O/r2c_stream_s$s.ncu-rep --page raw --csv > $O/r2c_stream_s$s.csv 2>/dev/null
done
python tools/ncu_summary.py $O/r2c_stream_s4.csv $O/r2c_stream_s8.csv $O/r2c_stream_s16.csv > $O/r2c_ncu_summary.csv
cat $O/r2c_ncu_summary.csv | cut -c1-1500
