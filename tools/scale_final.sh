# Short 8-GPU confirmation run of the final build (gpurun --gpus 8): C-ABI multi-GPU tests, N = 8 of the headline (sample-sharded) and of
# config 5 (tile-sharded, 8-row bands, box / sinc / gaussian on the gathered image), each with its N = 1 baseline on the same box.
TAG=${TAG:-final}
timeout 300 python -m pytest tests/test_gpu_multi.py -x -q -m gpu 2>&1 | tail -2
run() {  # workload sharding n
  local out=gpurun_out/${TAG}_$1_$2_$3
  if [ $3 = 1 ]; then timeout 200 python bench.py --gpus 1 --steps ${STEPS:-40} --warmup 5 --no-cpu --workload $1 --sharding $2 > $out.json 2> $out.err
  else timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $3 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $3 --steps ${STEPS:-40} --warmup 5 --no-cpu --workload $1 --sharding $2 > $out.json 2> $out.err; fi
  python - $out $1 $2 $3 <<'PY'
import json, sys
out, wl, sh, n = sys.argv[1:5]
try:
    d = json.loads(open(out + ".json").read().strip().splitlines()[-1])
    c = d.get("combine") or {}
    print("%-8s %-6s N=%s value %9.1f  e2e %9.1f (sync %9.1f)  ms/step %.3f  combine %s ms" % (wl, sh, n, d["value"], d["e2e"]["value"], d["e2e"]["synchronous_value"], d["ms_per_step"], "%.3f" % c["ms"] if c.get("ms") else "-"))
except Exception as e:
    print("%s %s N=%s failed: %s" % (wl, sh, n, e)); print(open(out + ".err").read()[-600:])
PY
}
run cornell sample 1; run cornell sample 8
run config5 tile 1; run config5 tile 8
