#!/bin/bash
# round 2, final build: smoke, whole GPU suite, full bench + reference arm, launch list, ncu --set full of the headline
# kernels and of k_warp2, traffic.json
O=gpurun_out
mkdir -p $O
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/r2z_smoke.log 2>&1; tail -2 $O/r2z_smoke.log
timeout 900 python -m pytest tests -m gpu -x -q > $O/r2z_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/r2z_pytest_gpu.log; tail -3 $O/r2z_pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 > $O/r2z_bench_full.json 2> $O/r2z_bench_full.err; python tools/show_bench.py $O/r2z_bench_full.json; tail -2 $O/r2z_bench_full.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > $O/r2z_bench_reference.json 2> $O/r2z_bench_reference.err; cut -c1-200 $O/r2z_bench_reference.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2z_launches_bench_tree_d10_3x3x1.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-extra > $O/r2z_ncu_launch.log 2>&1
python tools/one_call.py A 0 > $O/r2z_one.log 2>&1 && ncu --set full --import-source on --clock-control none -k regex:"k_stream2|k_jump_apply" -c 4 -o $O/r2z_stream -f python tools/one_call.py A 0 > $O/r2z_ncu_full.log 2>&1
ncu -i $O/r2z_stream.ncu-rep --page raw --csv > $O/r2z_stream_raw.csv 2>/dev/null
ncu --set full --import-source on --clock-control none -k regex:"k_warp2" -c 1 -o $O/r2z_warp2 -f python tools/one_call_dr.py 20 1 3 1 1 > $O/r2z_ncu_warp2.log 2>&1
ncu -i $O/r2z_warp2.ncu-rep --page raw --csv > $O/r2z_warp2_raw.csv 2>/dev/null
python tools/ncu_summary.py $O/r2z_stream_raw.csv $O/r2z_warp2_raw.csv > $O/r2z_ncu_full_summary.csv; cut -c1-300 $O/r2z_ncu_full_summary.csv | tail -5
python tools/make_traffic.py tree_d10_3x3x1 k_stream2 $O/r2z_stream_raw.csv profiles/r2z_ncu_full_summary.csv $O/traffic.json
python tools/make_traffic.py tree_d20_9x9x3_dr k_warp2 $O/r2z_warp2_raw.csv profiles/r2z_ncu_full_summary.csv $O/traffic.json
rm -f $O/r2z_warp2.ncu-rep
