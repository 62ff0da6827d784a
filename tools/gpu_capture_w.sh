# usage: TAG=... W=config2 K=mega bash tools/gpu_capture_w.sh   -- one ncu --set full capture of a non-default workload
set -x
python bench.py --steps 2 --warmup 3 --no-cpu --workload $W --kernel $K > gpurun_out/plain_${TAG}.log 2>&1 || exit 1
SAIL_B200_JIT_DUMP=gpurun_out ncu --set full --clock-control none --import-source on -k regex:sail_jit_kernel\|sail_trace -s 3 -c 1 -f -o gpurun_out/prof_${TAG} python bench.py --steps 2 --warmup 3 --no-cpu --workload $W --kernel $K > gpurun_out/ncu_${TAG}.log 2>&1
