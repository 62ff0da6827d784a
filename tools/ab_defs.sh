# A/B of experimental JIT code paths: tools/ab_defs.sh "DEFS1" "DEFS2" ... (each a SAIL_B200_JIT_DEFINES value; "-" = none)
for d in "$@"; do
  if [ "$d" = "-" ]; then unset SAIL_B200_JIT_DEFINES; else export SAIL_B200_JIT_DEFINES="$d"; fi
  for wl in ${WORKLOADS:-cornell}; do
    timeout 200 python bench.py --workload $wl --kernel ${KERNEL:-auto} --steps ${STEPS:-30} --warmup 3 --no-cpu > gpurun_out/ab.json 2> gpurun_out/ab.err || { echo "defs=$d wl=$wl FAILED"; tail -3 gpurun_out/ab.err; continue; }
    python - "$d" "$wl" <<'PY'
import json, sys
d = json.load(open("gpurun_out/ab.json"))
print("defs=%-40s wl=%-10s %8.1f Mpaths/s frac %.4f  %s" % (sys.argv[1], sys.argv[2], d["value"], d["roofline"]["frac"], d["config"]["kernel_used"][:60]))
PY
  done
done
