for w in cornell config2 config3 config5; do for k in mega wavefront; do for j in "" "--no-jit"; do
python bench.py --steps 6 --warmup 2 --no-cpu --kernel $k --workload $w $j | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('%-8s %-10s %9.1f Mpaths/s  frac %.3f  %s' % ('$w', d['config']['kernel'], d['value'], d['roofline']['frac'], d['config']['kernel_used'][:70]))"
done; done; done
