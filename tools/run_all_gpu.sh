set -x
timeout 1200 python -m pytest tests -x -q -m gpu 2>&1 | tail -5 > gpurun_out/r1_gpu_tests.log; cat gpurun_out/r1_gpu_tests.log
timeout 300 python __graft_entry__.py --smoke 2>&1 | tail -2
timeout 600 python bench.py > gpurun_out/b_c4.json 2> gpurun_out/b_c4.err; echo rc $?
timeout 600 python bench.py --impl reference > gpurun_out/b_c4_ref.json 2> gpurun_out/b_c4_ref.err; echo rc $?
for w in c1 c3; do timeout 300 python bench.py --workload $w > gpurun_out/b_$w.json 2> gpurun_out/b_$w.err; echo rc $?; done
timeout 600 python bench.py --workload c5 --c5-mode local > gpurun_out/b_c5.json 2> gpurun_out/b_c5.err; echo rc $?
timeout 600 python bench.py --workload c5 --c5-mode random > gpurun_out/b_c5r.json 2> gpurun_out/b_c5r.err; echo rc $?
timeout 600 python bench.py --workload c5 --c5-mode random --solver dense --no-cpu > gpurun_out/b_c5r_dense.json 2> gpurun_out/b_c5r_dense.err; echo rc $?
B2BA_DENSE_MODE=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:dense_update --launch-skip 1 --launch-count 1 -o gpurun_out/r1_dense_update python bench.py --workload c5 --c5-mode random --solver dense --steps 1 --warmup 0 --iters 1 --no-cpu > gpurun_out/ncu_dense_full.log 2>&1; echo rc $?
ls -la gpurun_out | tail -20
