"""Generates tests/golden/oracle_v1.npz from the CPU oracle (oracle/oracle.cpp).

The reference ships no golden vectors and cannot run here (GLSL + JS, no graphics
stack, no Node), so these fixtures pin the ORACLE: any later change to oracle/ or detmath.h that alters a
bit shows up as a test failure, and the CUDA path is compared with the same files on the GPU box.
Run: python tests/golden/make_golden.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle as oracle_pkg  # noqa: E402
from sail_b200 import configs, schedule, wire  # noqa: E402

SIZES = {1: (32, 32), 2: (48, 27), 3: (48, 27), 5: (48, 27)}
FRAMES = 2


def main():
    oracle_pkg.build()
    orc = oracle_pkg.Oracle()
    out = {}
    for cfg, (W, H) in SIZES.items():
        scene, _, _, _, bounces = configs.CONFIGS[cfg](W, H)
        o, t, l = wire.serialize_scene(scene)
        orc.set_scene(o, t, l)
        sch = schedule.FrameSchedule()
        acc = np.zeros((H, W, 3), np.float32)
        for k in range(FRAMES):
            m = schedule.frame_matrix(scene.mat, sch.jitter_unit(k), W, H)
            r = orc.render_frame(W, H, m, scene.eye, k / (k + 1.0), sch.time(k), acc, max_bounces=bounces, aovs=True)
        out["c%d_objects" % cfg], out["c%d_texparams" % cfg], out["c%d_lights" % cfg] = o, t, l
        out["c%d_color" % cfg], out["c%d_hit_id" % cfg], out["c%d_depth" % cfg] = acc, r["hit_id"], r["depth"]
        out["c%d_normal" % cfg] = r["normal"]
    xs = np.float32([0.0, 0.5, -1.25, 3.14159274, 100.0, 12345.678, -250000.0, 1e-6])
    for name in ("sin", "cos", "tan", "atan", "exp"):
        out["dm_" + name] = orc.detmath(name, xs)
    us = np.float32([-1.0, -0.75, -0.1, 0.0, 0.3, 0.5, 0.99, 1.0])
    out["dm_acos"], out["dm_asin"] = orc.detmath("acos", us), orc.detmath("asin", us)
    out["dm_log"] = orc.detmath("log", np.abs(xs) + np.float32(1e-3))
    out["dm_x"], out["dm_u"] = xs, us
    out["rng"] = np.stack([orc.random2(x, y, s) for (x, y, s) in [(0, 0, 1.0), (17, 5, 2.0166667), (511, 511, 5.5), (1919, 1079, 3.25)]])
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "oracle_v1.npz"), **out)
    print("wrote oracle_v1.npz:", {k: v.shape for k, v in out.items() if k.startswith("c1")})


if __name__ == "__main__":
    main()
