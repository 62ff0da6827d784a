set -x
python bench.py > gpurun_out/bench_${TAG:-r1x}.json 2> gpurun_out/bench_${TAG:-r1x}.err || exit 1
cat gpurun_out/bench_${TAG:-r1x}.json
python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/plain_${TAG:-r1x}.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_${TAG:-r1x}.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_l.log 2>&1
SAIL_B200_JIT_DUMP=gpurun_out ncu --set full --clock-control none --import-source on -k regex:sail_jit_kernel\|megakernel -s 3 -c 1 -f -o gpurun_out/prof_${TAG:-r1x} python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_f.log 2>&1
ls -la gpurun_out
