#!/bin/bash
O=gpurun_out
TAG=${1:-r2o}
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "warp2 or extreme or general_path or lane_boundary or words_mode or private" > $O/${TAG}_pytest_warp.log 2>&1; echo "pytest rc=$?" >> $O/${TAG}_pytest_warp.log; tail -3 $O/${TAG}_pytest_warp.log
python tools/one_call_dr.py 20 0 > $O/${TAG}_one.log 2>&1
for lock in 1 2 4; do python tools/one_call_dr.py 20 1 3 $lock >> $O/${TAG}_one.log 2>&1; done
for lock in 1 2 4; do python tools/one_call_dr.py 20 1 1 $lock >> $O/${TAG}_one.log 2>&1; done
cat $O/${TAG}_one.log
for lock in 1 2 4; do
ncu --set full --import-source on --clock-control none -k regex:"k_warp2" -c 1 -o $O/${TAG}_warp2_l$lock -f python tools/one_call_dr.py 17 1 3 $lock > $O/${TAG}_ncu_l$lock.log 2>&1
ncu -i $O/${TAG}_warp2_l$lock.ncu-rep --page raw --csv > $O/${TAG}_warp2_l${lock}_raw.csv 2>/dev/null
done
python tools/ncu_summary.py $O/${TAG}_warp2_l1_raw.csv $O/${TAG}_warp2_l2_raw.csv $O/${TAG}_warp2_l4_raw.csv > $O/${TAG}_ncu_summary.csv; tail -3 $O/${TAG}_ncu_summary.csv | cut -c1-400
ncu -i $O/${TAG}_warp2_l1.ncu-rep --page source --csv > $O/${TAG}_warp2_l1_source.csv 2>/dev/null
rm -f $O/${TAG}_warp2_l*.ncu-rep
