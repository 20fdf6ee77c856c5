#!/bin/bash
O=gpurun_out
ncu --set full --import-source on --clock-control none -k regex:"k_warp2" -c 1 -o $O/r2u_warp2 -f python tools/one_call_dr.py 17 1 3 1 1 > $O/r2u_ncu.log 2>&1
ncu -i $O/r2u_warp2.ncu-rep --page raw --csv > $O/r2u_warp2_raw.csv 2>/dev/null
ncu -i $O/r2u_warp2.ncu-rep --page source --csv > $O/r2u_warp2_source.csv 2>/dev/null
rm -f $O/r2u_warp2.ncu-rep
python tools/ncu_summary.py $O/r2u_warp2_raw.csv | tail -1 | cut -c1-300
