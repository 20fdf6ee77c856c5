#!/bin/bash
# round 2: evidence run — sanitizer logs, launch list, ncu --set full of the headline kernels, full bench, reference arm
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "geometry or packed or zero_copy or full_size_properties or multi" > $O/r2j_pytest.log 2>&1; echo "pytest rc=$?" >> $O/r2j_pytest.log; tail -3 $O/r2j_pytest.log
python tools/sanitize_small.py > $O/r2j_sanitize_plain.log 2>&1; tail -2 $O/r2j_sanitize_plain.log
timeout 900 compute-sanitizer --tool memcheck --print-limit 20 python tools/sanitize_small.py > $O/r2j_sanitizer_memcheck.log 2>&1; tail -3 $O/r2j_sanitizer_memcheck.log
timeout 1200 compute-sanitizer --tool racecheck --print-limit 20 python tools/sanitize_small.py > $O/r2j_sanitizer_racecheck.log 2>&1; tail -3 $O/r2j_sanitizer_racecheck.log
timeout 600 python bench.py --steps 10 --warmup 3 > $O/r2j_bench_full.json 2> $O/r2j_bench_full.err; python tools/show_bench.py $O/r2j_bench_full.json
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > $O/r2j_bench_reference.json 2> $O/r2j_bench_reference.err; cut -c1-400 $O/r2j_bench_reference.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2j_launches_bench_tree_d10_3x3x1.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-extra > $O/r2j_ncu_launch.log 2>&1
python tools/one_call.py A 0 > $O/r2j_one.log 2>&1 && ncu --set full --import-source on --clock-control none -k regex:"k_stream2|k_jump_apply" -c 4 -o $O/r2j_stream -f python tools/one_call.py A 0 > $O/r2j_ncu_full.log 2>&1
ncu -i $O/r2j_stream.ncu-rep --page raw --csv > $O/r2j_stream_raw.csv 2>/dev/null
python tools/ncu_summary.py $O/r2j_stream_raw.csv > $O/r2j_ncu_full_summary.csv; cut -c1-900 $O/r2j_ncu_full_summary.csv | tail -4
python tools/make_traffic.py tree_d10_3x3x1 k_stream2 $O/r2j_stream_raw.csv profiles/r2j_ncu_full_summary.csv $O/traffic.json
