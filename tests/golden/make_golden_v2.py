"""Generates tests/golden/oracle_v2.npz: the long-accumulation and per-bounce fixtures SURVEY.md 8(c)(2) asks for.

  c1_converged : config 1 (README Cornell box) 64 x 64, 256 progressive samples (the running mean of fstrace.glsl:14)
  c2_converged : config 2 (materials demo)     48 x 27, 128 progressive samples
  c2_bounce_*  : frame 0 of config 2, per-bounce hit id / hit distance / ray (origin, direction) of every path
Same purpose as oracle_v1.npz (pins the oracle; the CUDA path is compared with the same file on the GPU box).
Run: python tests/golden/make_golden_v2.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle as oracle_pkg  # noqa: E402
from sail_b200 import configs, schedule, wire  # noqa: E402

CONVERGED = {1: (64, 64, 256), 2: (48, 27, 128)}


def main():
    oracle_pkg.build()
    orc = oracle_pkg.Oracle()
    out = {}
    for cfg, (W, H, spp) in CONVERGED.items():
        scene, _, _, _, bounces = configs.CONFIGS[cfg](W, H)
        orc.set_scene(*wire.serialize_scene(scene))
        sch = schedule.FrameSchedule()
        acc = np.zeros((H, W, 3), np.float32)
        for k in range(spp):
            m = schedule.frame_matrix(scene.mat, sch.jitter_unit(k), W, H)
            orc.render_frame(W, H, m, scene.eye, k / (k + 1.0), sch.time(k), acc, max_bounces=bounces)
        out["c%d_converged" % cfg] = acc
    W, H, _ = CONVERGED[2]
    scene, _, _, _, bounces = configs.CONFIGS[2](W, H)
    orc.set_scene(*wire.serialize_scene(scene))
    sch = schedule.FrameSchedule()
    m = schedule.frame_matrix(scene.mat, sch.jitter_unit(0), W, H)
    d = orc.trace_dump(W, H, m, scene.eye, sch.time(0), max_bounces=bounces)
    for k, v in d.items():
        out["c2_bounce_" + k] = v
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "oracle_v2.npz"), **out)
    print({k: (v.shape, str(v.dtype)) for k, v in out.items()})


if __name__ == "__main__":
    main()
