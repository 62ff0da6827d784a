# 1 -> 8 GPU runs of bench.py on one box (gpurun --gpus 8): one JSON line per run into gpurun_out/<TAG>_<workload>_<sharding>[_bB]_N.json
#   sample sharding (weak scaling): cornell (headline) and config4 (BASELINE config 4: 4K Cornell);
#   tile sharding (strong scaling): config5 (BASELINE config 5: glass + metal, 16 bounces, filter sweep box / sinc / gaussian
#   on the gathered image), band heights 8 / 16 / 32.
# The combine of the GPUs' buffers and the present of the COMBINED image are inside the timed e2e region (bench.py).
TAG=${TAG:-scale}
run() {  # workload sharding band n
  local wl=$1 sh=$2 band=$3 n=$4 out=gpurun_out/${TAG}_$1_$2$([ $3 != 0 ] && echo _b$3)_$4
  if [ $n = 1 ]; then timeout 300 python bench.py --gpus 1 --steps ${STEPS:-40} --warmup 5 --no-cpu --workload $wl --sharding $sh > $out.json 2> $out.err
  else timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $n --steps ${STEPS:-40} --warmup 5 --no-cpu --workload $wl --sharding $sh --band-rows $band > $out.json 2> $out.err; fi
  python - $out $wl $sh $band $n <<'PY'
import json, sys
out, wl, sh, band, n = sys.argv[1:6]
try:
    d = json.loads(open(out + ".json").read().strip().splitlines()[-1])
    c = d.get("combine") or {}
    print("%-8s %-6s band %-2s N=%s value %9.1f Mpaths/s  e2e %9.1f (sync %9.1f)  ms/step %.3f  trace min/max %s/%s  combine %s ms %s GB/s" % (
        wl, sh, band, n, d["value"], d["e2e"]["value"], d["e2e"]["synchronous_value"], d["ms_per_step"],
        "%.3f" % d["trace_ms_per_step_min_over_ranks"] if "trace_ms_per_step_min_over_ranks" in d else "-",
        "%.3f" % d["trace_ms_per_step_max_over_ranks"] if "trace_ms_per_step_max_over_ranks" in d else "-",
        "%.3f" % c["ms"] if c.get("ms") else "-", "%.0f" % c["gbs"] if c.get("gbs") else "-"))
except Exception as e:
    print("%s %s N=%s failed: %s" % (wl, sh, n, e)); print(open(out + ".err").read()[-800:])
PY
}
for n in ${NS:-1 2 4 8}; do run cornell sample 0 $n; done
for n in ${NS:-1 2 4 8}; do run config4 sample 0 $n; done
run config5 tile 0 1
for b in 8 16 32; do for n in ${NS:-1 2 4 8}; do [ $n = 1 ] || run config5 tile $b $n; done; done
