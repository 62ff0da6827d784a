"""Device-free NVRTC build of a BASELINE config's kernel: python tools/jit_check.py <config 1|2|3|5> <kernel 0 mega|1 wavefront>
(SAIL_B200_JIT_DUMP=<dir> keeps the cubin; SAIL_B200_JIT_DEFINES adds #defines)."""
import os
import sys
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from sail_b200 import capi
from conftest import scene_tables

cfg, kern = int(sys.argv[1]), int(sys.argv[2])
scene, o, t, l, b = scene_tables(cfg, 64, 64)
print(capi.jit_compile_check(o, t, l, kernel=kern))
