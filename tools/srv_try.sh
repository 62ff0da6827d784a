# Server-warp kernel: parity first (bounded: a protocol bug would hang), then the CTA-shape sweep on the 1080p Cornell box.
set -x
timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "server or tile_rows or interleaved" > gpurun_out/srv_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/srv_pytest.log
for w in 3 5 7 11 15; do
  SAIL_B200_SRV_WARPS=$w timeout 120 python bench.py --kernel server --steps 20 --warmup 3 --no-cpu > gpurun_out/srv_w$w.json 2> gpurun_out/srv_w$w.err; echo "w=$w rc=$?"
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/srv_w$w.json")); print("w=$w", round(d["value"]), d["roofline"]["frac"], d["config"]["kernel_used"])
except Exception as e: print("w=$w failed", e)
PY
done
timeout 120 python bench.py --kernel mega --steps 20 --warmup 3 --no-cpu > gpurun_out/srv_mega.json 2>&1; python -c "
import json; d=json.load(open('gpurun_out/srv_mega.json')); print('mega', round(d['value']), d['roofline']['frac'])"
