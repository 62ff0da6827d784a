#!/usr/bin/env python
"""One-off structure fuzz beyond the suite's ten seeds: python tools/fuzz_scenes.py FIRST COUNT  (CUDA jit + aot vs oracle, bit for bit)."""
import os
import sys

import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
os.environ.setdefault("SAIL_B200_JIT_SYNC", "1")
import matrix_scenes  # noqa: E402
import oracle as oracle_pkg  # noqa: E402
from conftest import bits_equal, frame_uniforms  # noqa: E402
from sail_b200 import capi, wire  # noqa: E402

first, count = int(sys.argv[1]), int(sys.argv[2])
orc = oracle_pkg.Oracle()
W, H = 64, 48
bad = 0
for seed in range(first, first + count):
    scene = matrix_scenes.random_scene(seed, W, H)
    o, t, l = wire.serialize_scene(scene)
    orc.set_scene(o, t, l)
    acc = np.zeros((H, W, 3), np.float32)
    frames = []
    for k in range(2):
        m, eye, w, tm = frame_uniforms(scene, k, W, H)
        ref = orc.render_frame(W, H, m, eye, w, tm, acc, max_bounces=5, aovs=True)
        frames.append(capi.make_frame(m, eye, w, tm))
    for kernel, spec in ((seed % 2, 2), (1 - seed % 2, 1), (1 - seed % 2, 2)):
        with capi.Context(W, H) as ctx:
            ctx.configure(kernel=kernel, specialise=spec)
            ctx.upload_scene(o, t, l)
            ctx.render_n(frames)
            ok = bits_equal(ctx.read_accum(), acc) and np.array_equal(ctx.read_aov(capi.AOV_HIT_ID), ref["hit_id"]) and \
                bits_equal(ctx.read_aov(capi.AOV_DEPTH), ref["depth"]) and bits_equal(ctx.read_aov(capi.AOV_NORMAL), ref["normal"])
            if not ok:
                bad += 1
                print("MISMATCH seed %d kernel %d specialise %d: %s" % (seed, kernel, spec, ctx.kernel_info()), flush=True)
print("fuzz: seeds %d..%d, %d mismatches" % (first, first + count - 1, bad))
sys.exit(1 if bad else 0)
