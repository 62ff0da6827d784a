# usage: TAG=r02c WORKLOADS="cornell config3" bash tools/gpu_capture2.sh
# Per workload: a plain run (must exit 0), the ncu launch list of the same command, one full capture of the trace kernel with source
# counters, the dumped cubin for tools/sass_lines.py.  Run under gpurun on ONE GPU; bring gpurun_out/ back and summarise into profiles/.
set -x
for wl in ${WORKLOADS:-cornell}; do
  T=${TAG:-r02}_$wl
  python bench.py --workload $wl --steps 2 --warmup 3 --no-cpu > gpurun_out/plain_$T.json 2> gpurun_out/plain_$T.err || { tail -5 gpurun_out/plain_$T.err; continue; }
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$T.csv python bench.py --workload $wl --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_l_$T.log 2>&1
  mkdir -p gpurun_out/cubin_$T
  SAIL_B200_JIT_DUMP=gpurun_out/cubin_$T ncu --set full --clock-control none --import-source on -k regex:sail_jit_kernel -s 3 -c 1 -f -o gpurun_out/prof_$T python bench.py --workload $wl --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_f_$T.log 2>&1
done
ls -la gpurun_out | tail -20
