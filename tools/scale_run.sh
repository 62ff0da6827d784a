# 1 -> 8 GPU scaling of bench.py on one box (gpurun --gpus 8): one JSON line per N into gpurun_out/scale_N.json
for n in ${NS:-1 2 4 8}; do
  if [ $n = 1 ]; then python bench.py --gpus 1 --steps ${STEPS:-40} --warmup 5 --no-cpu --workload ${WORKLOAD:-cornell} > gpurun_out/scale_${WORKLOAD:-cornell}_$n.json 2> gpurun_out/scale_${WORKLOAD:-cornell}_$n.err
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $n --steps ${STEPS:-40} --warmup 5 --no-cpu --workload ${WORKLOAD:-cornell} > gpurun_out/scale_${WORKLOAD:-cornell}_$n.json 2> gpurun_out/scale_${WORKLOAD:-cornell}_$n.err; fi
  python - $n ${WORKLOAD:-cornell} <<'PY'
import json, sys
n = sys.argv[1]
try:
    d = json.loads(open("gpurun_out/scale_%s_%s.json" % (sys.argv[2], n)).read().strip().splitlines()[-1])
    print("N=%s value %.1f Mpaths/s  e2e %.1f  ms/step %.3f  %s" % (n, d["value"], d["e2e"]["value"], d["ms_per_step"], d.get("reduce_ms", "")))
except Exception as e:
    print("N=%s failed: %s" % (n, e)); print(open("gpurun_out/scale_%s_%s.err" % (sys.argv[2], n)).read()[-800:])
PY
done
[ -n "$SKIP_MULTI_TEST" ] || timeout 300 python -m pytest tests/test_gpu_multi.py -x -q -m gpu 2>&1 | tail -2
