# multi-rank pass on one box with 8 GPUs: correctness at 2 / 4 / 8 ranks, C5 strong scaling at 1 / 2 / 4 / 8, bench line at 8
set -x
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for n in 2 4 8; do
  timeout 240 $TR --nproc-per-node $n --master-port 2950$n tools/multi_gpu_check.py > gpurun_out/r2_multi_gpu_check_${n}ranks.log 2>&1; echo rc $?
  tail -6 gpurun_out/r2_multi_gpu_check_${n}ranks.log
done
timeout 200 python tools/c5_scaling.py > gpurun_out/r2_c5_scaling_1gpu.json 2> gpurun_out/c5s1.err; echo rc $?
for n in 2 4 8; do
  timeout 240 $TR --nproc-per-node $n --master-port 2951$n tools/c5_scaling.py > gpurun_out/r2_c5_scaling_${n}gpu.json 2> gpurun_out/c5s$n.err; echo rc $?
done
cat gpurun_out/r2_c5_scaling_*gpu.json
timeout 300 $TR --nproc-per-node 8 --master-port 29528 bench.py --gpus 8 --no-cpu > gpurun_out/r2_bench_c4_8gpu.json 2> gpurun_out/b8.err; echo rc $?
