# Wavefront CTA size sweep (SAIL_B200_WF_THREADS) with a parity check per size
for w in 256 512; do
  SAIL_B200_WF_THREADS=$w timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "frames_bit_exact and wavefront and jit or short_paths and wavefront or interleaved or tile_rows" 2>&1 | tail -1
done
for w in 128 256 512; do
  export SAIL_B200_WF_THREADS=$w
  echo "== wavefront paths per CTA: $w"
  KERNEL=wavefront WORKLOADS="${WORKLOADS:-config3 config2 config5}" STEPS=8 bash tools/ab_defs.sh -
done
