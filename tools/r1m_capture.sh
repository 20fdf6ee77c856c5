set -u
O=gpurun_out
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/r1m_smoke.log 2>&1; tail -1 $O/r1m_smoke.log
python bench.py --steps 10 --warmup 3 > $O/r1m_bench_tree_d10_3x3x1.json 2> $O/r1m.err
for w in tree_d20_9x9x3 mold_2_10 tree_d10_3x3x1_dr tree_d20_9x9x3_dr; do python bench.py --steps 10 --warmup 3 --no-cpu --workload $w > $O/r1m_bench_$w.json 2>> $O/r1m.err; done
python bench.py --steps 20 --warmup 3 --workload search_offpolicy > $O/r1m_bench_search.json 2>> $O/r1m.err
python bench.py --impl reference --steps 3 --warmup 1 > $O/r1m_bench_reference.json 2>> $O/r1m.err
python tools/show_bench.py $O/r1m_bench_*.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r1m_launches_bench_tree_d10_3x3x1.csv python bench.py --steps 2 --warmup 3 --no-cpu > $O/r1m_ncu_launch.log 2>&1
python tools/one_call.py A 4 > /dev/null 2>&1 && ncu --set full --import-source on --clock-control none -k regex:"k_stream2|k_jump_apply" -c 4 -o $O/r1m_stream -f python tools/one_call.py A 4 > $O/r1m_ncu_full.log 2>&1
ls -la $O/r1m_stream.ncu-rep
