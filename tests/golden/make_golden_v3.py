"""Writes tests/golden/oracle_v3.npz: fixtures of the CPU oracle (oracle/) for what round 2 added -- the plugin-matrix scenes
(tests/matrix_scenes.py: 3-frame running mean + primary hit ids at 48x32), the opt-in Russian roulette, and SHA-256 digests of
one full frame (k = 2) of every BASELINE config AT ITS STATED SIZE (512x512, 1920x1080, 3840x2160: colour, hit ids, depth), so
the GPU suite can check full-size frames with no oracle in the loop.  PARITY UNPINNED against the reference itself (it holds
no vectors and cannot run here); this pins the CUDA path to the oracle as committed.

    python tests/golden/make_golden_v3.py
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle as oracle_pkg  # noqa: E402
import matrix_scenes  # noqa: E402
from conftest import frame_uniforms, scene_tables  # noqa: E402
from sail_b200 import wire  # noqa: E402

FULL_SIZES = {1: (512, 512), 2: (1920, 1080), 3: (1920, 1080), 4: (3840, 2160), 5: (1920, 1080)}


def digest(a):
    a = np.ascontiguousarray(a)
    if a.dtype == np.float32:
        bits = a.view(np.uint32).copy()
        bits[np.isnan(a)] = 0x7FC00000
        a = bits
    return np.frombuffer(hashlib.sha256(a.tobytes()).digest(), np.uint8)


def main():
    oracle_pkg.build()
    orc = oracle_pkg.Oracle()
    out = {}
    W, H = 48, 32
    for name, fn in sorted(matrix_scenes.MATRIX.items()):
        scene = fn(W, H)
        o, t, l = wire.serialize_scene(scene)
        orc.set_scene(o, t, l)
        acc = np.zeros((H, W, 3), np.float32)
        for k in range(3):
            m, eye, w, tm = frame_uniforms(scene, k, W, H)
            r = orc.render_frame(W, H, m, eye, w, tm, acc, max_bounces=5, aovs=True)
        out["m_%s_color" % name] = acc
        out["m_%s_hit_id" % name] = r["hit_id"].astype(np.int8)
    # Russian roulette from bounce 2 on config 2
    scene, o, t, l, bounces = scene_tables(2, W, H)
    orc.set_scene(o, t, l)
    orc.set_russian_roulette(2)
    acc = np.zeros((H, W, 3), np.float32)
    for k in range(4):
        m, eye, w, tm = frame_uniforms(scene, k, W, H)
        orc.render_frame(W, H, m, eye, w, tm, acc, max_bounces=bounces)
    orc.set_russian_roulette(0)
    out["rr2_c2_color"] = acc
    # full-size frames: digests only
    for cfg, (FW, FH) in FULL_SIZES.items():
        scene, o, t, l, bounces = scene_tables(cfg, FW, FH)
        orc.set_scene(o, t, l)
        acc = np.zeros((FH, FW, 3), np.float32)
        m, eye, w, tm = frame_uniforms(scene, 2, FW, FH)
        r = orc.render_frame(FW, FH, m, eye, w, tm, acc, max_bounces=bounces, aovs=True)
        out["full_c%d_color_sha256" % cfg] = digest(acc)
        out["full_c%d_hit_id_sha256" % cfg] = digest(r["hit_id"])
        out["full_c%d_depth_sha256" % cfg] = digest(r["depth"])
        print("config %d at %dx%d: mean %.6f" % (cfg, FW, FH, float(acc.mean())))
    np.savez_compressed(os.path.join(HERE, "oracle_v3.npz"), **out)
    print("wrote oracle_v3.npz:", os.path.getsize(os.path.join(HERE, "oracle_v3.npz")), "bytes,", len(out), "arrays")


if __name__ == "__main__":
    main()
