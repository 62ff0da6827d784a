# measures what FMA contraction would buy (parity is NOT expected to hold in this build)
set -e
python bench.py --steps 30 --warmup 3 --no-cpu | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('strict', d['value'])"
sed -i 's/-fmad=false/-fmad=true/' sail_b200/csrc/Makefile
make -C sail_b200/csrc clean >/dev/null; make -C sail_b200/csrc -j16 >/dev/null 2>&1
python bench.py --steps 30 --warmup 3 --no-cpu | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('fmad', d['value'])"
