set -e
for m in 5 6 7 8; do
  make -C sail_b200/csrc clean >/dev/null; make -C sail_b200/csrc -j16 NVCC="/usr/local/cuda/bin/nvcc -DSAIL_MIN_CTAS_SMALL=$m" >/dev/null 2>&1
  echo "MIN_CTAS=$m"; python bench.py --steps 20 --warmup 3 --no-cpu | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['roofline']['frac'])"
done
