"""Counts OPS_alg: FP32 arithmetic operations the literal oracle executes per path,
averaged over the frame and over frames k = 0..F-1, with the op-counting oracle build
(oracle/liboracle_ops.so, -DSAIL_COUNT_OPS).  Writes baseline/ops_per_path.json.
Every config over frames k = 0..15, as SURVEY.md 8(d) defines it.
Usage: python baseline/ops_per_path.py [frames]"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oracle as oracle_pkg  # noqa: E402
from sail_b200 import configs, schedule, wire  # noqa: E402


def count(builder, frames):
    scene, W, H, _, bounces = builder()
    o, t, l = wire.serialize_scene(scene)
    orc = oracle_pkg.Oracle(count_ops=True)
    orc.set_scene(o, t, l)
    sch = schedule.FrameSchedule()
    acc = np.zeros((H, W, 3), np.float32)
    total = 0
    for k in range(frames):
        m = schedule.frame_matrix(scene.mat, sch.jitter_unit(k), W, H)
        total += orc.render_frame(W, H, m, scene.eye, k / (k + 1.0), sch.time(k), acc, max_bounces=bounces)["ops"]
    return {"width": W, "height": H, "max_bounces": bounces, "frames": frames, "ops_per_path": total / (frames * W * H)}


if __name__ == "__main__":
    frames = int(sys.argv[1]) if len(sys.argv) > 1 else 16
    oracle_pkg.build()
    out = {"definition": "FP32 ops per path counted by the literal oracle: + - * / min max abs floor compare sqrt "
                         "and each transcendental = 1; addressing, conversions and loads = 0 (SURVEY.md 8(d))",
           "cornell_1080p": count(configs.cornell_1080p, frames),
           "config1_512": count(configs.config1, frames),
           "config2_1080p": count(configs.config2, frames),
           "config3_1080p": count(configs.config3, frames),
           "config5_1080p": count(configs.config5, frames),
           "config4_2160p": count(configs.config4, frames)}
    with open(os.path.join(ROOT, "baseline", "ops_per_path.json"), "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out, indent=1))
