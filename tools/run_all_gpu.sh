# one pass over everything the round reports: GPU tests, smoke, every bench workload, the ncu captures
set -x
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -5 > gpurun_out/r1_gpu_tests.log; cat gpurun_out/r1_gpu_tests.log
timeout 300 python __graft_entry__.py --smoke 2>&1 | tail -2
timeout 600 python bench.py > gpurun_out/b_c4.json 2> gpurun_out/b_c4.err; echo rc $?
timeout 600 python bench.py --impl reference > gpurun_out/b_c4_ref.json 2> gpurun_out/b_c4_ref.err; echo rc $?
timeout 900 python bench.py --workload c2 --frames 1000 > gpurun_out/b_c2.json 2> gpurun_out/b_c2.err; echo rc $?
timeout 900 python bench.py --workload c2 --frames 1000 --impl reference > gpurun_out/b_c2_ref.json 2> gpurun_out/b_c2_ref.err; echo rc $?
timeout 600 python bench.py --workload c5 --c5-mode random --solver dense --no-cpu > gpurun_out/b_c5r_dense.json 2> gpurun_out/b_c5r_dense.err; echo rc $?
B2BA_DENSE_MODE=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/r1_launches_c5r_dense.csv python bench.py --workload c5 --c5-mode random --solver dense --steps 1 --warmup 0 --iters 1 --no-cpu > gpurun_out/ncu_dense.log 2>&1; echo rc $?
ls -la gpurun_out | tail -12
for w in c1 c3; do timeout 300 python bench.py --workload $w > gpurun_out/b_$w.json 2> gpurun_out/b_$w.err; echo rc $?; done
timeout 600 python bench.py --workload c5 --c5-mode local > gpurun_out/b_c5.json 2> gpurun_out/b_c5.err; echo rc $?
timeout 600 python bench.py --workload c5 --c5-mode random > gpurun_out/b_c5r.json 2> gpurun_out/b_c5r.err; echo rc $?
timeout 300 python bench.py --workload pnp > gpurun_out/b_pnp.json 2> gpurun_out/b_pnp.err; echo rc $?
timeout 300 python tools/c2_profile.py 600 > gpurun_out/r1_c2_profile.txt 2>&1; echo rc $?
