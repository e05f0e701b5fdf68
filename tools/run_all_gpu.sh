# one pass over everything the round reports (1 GPU): GPU tests, smoke, bench lines of every workload, launch list, ncu capture
set -x
R=${ROUND:-r2}
timeout 900 python -m pytest tests -x -q -m gpu --durations=3 2>&1 | tail -8 > gpurun_out/${R}_gpu_tests.log; cat gpurun_out/${R}_gpu_tests.log
timeout 300 python __graft_entry__.py --smoke 2>&1 | tail -2
timeout 600 python bench.py > gpurun_out/${R}_bench_c4.json 2> gpurun_out/b_c4.err; echo rc $?
timeout 600 python bench.py --impl reference > gpurun_out/${R}_bench_c4_ref.json 2> gpurun_out/b_c4_ref.err; echo rc $?
for w in c1 c3; do timeout 300 python bench.py --workload $w > gpurun_out/${R}_bench_$w.json 2> gpurun_out/b_$w.err; echo rc $?; done
timeout 300 python bench.py --workload pnp > gpurun_out/${R}_bench_pnp.json 2> gpurun_out/b_pnp.err; echo rc $?
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${R}_launches_c4_4096win.csv python bench.py --steps 2 --warmup 1 --no-cpu --no-c5 > gpurun_out/ncu_l.log 2>&1; echo rc $?
timeout 500 ncu --set full --clock-control none --import-source on -k regex:window_lm -s 1 -c 1 -o gpurun_out/${R}_win592_final python bench.py --windows 592 --distinct 64 --steps 2 --warmup 1 --no-cpu --no-c5 > gpurun_out/ncu_w.log 2>&1; echo rc $?
if [ -z "$SKIP_C2" ]; then
timeout 900 python bench.py --workload c2 > gpurun_out/${R}_bench_c2.json 2> gpurun_out/b_c2.err; echo rc $?
timeout 900 python bench.py --workload c2 --impl reference > gpurun_out/${R}_bench_c2_ref.json 2> gpurun_out/b_c2_ref.err; echo rc $?
fi
ls -la gpurun_out | tail -20
