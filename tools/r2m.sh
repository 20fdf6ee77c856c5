#!/bin/bash
# k_warp2: parity of the warp path, old vs new timing, ncu --set full of the new kernel
O=gpurun_out
TAG=${1:-r2m}
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "warp2 or extreme or general_path or lane_boundary or words_mode or private" > $O/${TAG}_pytest_warp.log 2>&1; echo "pytest rc=$?" >> $O/${TAG}_pytest_warp.log; tail -4 $O/${TAG}_pytest_warp.log
python tools/one_call_dr.py 20 0 > $O/${TAG}_one.log 2>&1
python tools/one_call_dr.py 20 1 >> $O/${TAG}_one.log 2>&1
python tools/one_call_dr.py 20 1 1 >> $O/${TAG}_one.log 2>&1
cat $O/${TAG}_one.log
ncu --set full --import-source on --clock-control none -k regex:"k_warp2" -c 1 -o $O/${TAG}_warp2 -f python tools/one_call_dr.py 17 1 > $O/${TAG}_ncu.log 2>&1
ncu -i $O/${TAG}_warp2.ncu-rep --page raw --csv > $O/${TAG}_warp2_raw.csv 2>/dev/null
ncu -i $O/${TAG}_warp2.ncu-rep --page source --csv > $O/${TAG}_warp2_source.csv 2>/dev/null
python tools/ncu_summary.py $O/${TAG}_warp2_raw.csv > $O/${TAG}_ncu_summary.csv; tail -1 $O/${TAG}_ncu_summary.csv | cut -c1-600
