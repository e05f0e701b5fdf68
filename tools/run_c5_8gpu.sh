set -x
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 200 $TR --nproc-per-node 8 --master-port 29568 tools/multi_gpu_check.py > gpurun_out/r2_multi_gpu_check_8ranks.log 2>&1; echo rc $?; grep -c OK gpurun_out/r2_multi_gpu_check_8ranks.log
timeout 200 $TR --nproc-per-node 8 --master-port 29569 tools/c5_scaling.py 2>/dev/null | grep "^{" > gpurun_out/r2_c5_scaling_8gpu.json; echo rc $?
timeout 200 $TR --nproc-per-node 4 --master-port 29564 tools/c5_scaling.py 2>/dev/null | grep "^{" > gpurun_out/r2_c5_scaling_4gpu.json; echo rc $?
cat gpurun_out/r2_c5_scaling_8gpu.json gpurun_out/r2_c5_scaling_4gpu.json
