# Occupancy sweep of the big kernel variants (glossy / quadric scenes) on the GPU box: rebuilds with SAIL_MIN_CTAS_BIG=m.
for m in ${MS:-4 5 6}; do
  rm -f sail_b200/csrc/*.o; make -C sail_b200/csrc -j16 EXTRA=-DSAIL_MIN_CTAS_BIG=$m >/dev/null 2>&1 || { echo build failed; exit 1; }
  for w in config2 config3 config5; do for k in mega wavefront; do
    python bench.py --steps 6 --warmup 2 --no-cpu --kernel $k --workload $w | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('MIN_CTAS_BIG=$m %-8s %-10s %9.1f Mpaths/s' % ('$w', d['config']['kernel'], d['value']))"
  done; done
done
