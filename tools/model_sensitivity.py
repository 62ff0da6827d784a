#!/usr/bin/env python
"""How far apart are the two defensible readings of GLSL's '/' ?  ("What's weak" 1.)

The oracle -- and, bit for bit, the CUDA kernels -- define a / b := a * rcp(b) with a correctly rounded reciprocal and
normalize(v) := v * rcp(length(v)) (oracle/glsl.h).  SURVEY.md 8(c) first wrote "'/' correctly rounded".  Both lie inside
GLSL ES 3.00's 2.5 ulp allowance for division; the reference pins neither.  This tool renders the same frames with both
models (liboracle.so vs liboracle_ieee.so, built with -DSAIL_ORACLE_IEEE_DIV) and reports, per config:
  * primary and per-bounce hit-id agreement over frames 0..3 (oracle_trace_dump),
  * max-abs / mean-abs difference and PSNR of the N-spp running means, unclamped and clamped to [0, 1]
    (north_star: per-pixel max-abs error plus PSNR >= 40 dB against the converged image).

    python tools/model_sensitivity.py [--size 512] [--spp 1024] [--configs 1,2,3] [--out baseline/model_sensitivity.json]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle as oracle_pkg  # noqa: E402
from conftest import frame_uniforms, scene_tables  # noqa: E402


def psnr(a, b, peak):
    mse = float(np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2))
    return float("inf") if mse == 0 else 10.0 * np.log10(peak * peak / mse)


def compare(cfg, W, H, spp, threads=0, dump_frames=4):
    scene, o, t, l, bounces = scene_tables(cfg, W, H)
    models = {"rcp": oracle_pkg.Oracle(), "ieee": oracle_pkg.Oracle(ieee_div=True)}
    for m in models.values():
        m.set_scene(o, t, l)
    res = {"config": cfg, "size": [W, H], "spp": spp, "bounces": bounces}
    # ---- hit ids, bounce by bounce (both models trace their own path: after the first differing rounding the rays differ too)
    same = np.zeros(bounces, np.int64)
    total = np.zeros(bounces, np.int64)
    for k in range(dump_frames):
        mtx, eye, w, tm = frame_uniforms(scene, k, W, H)
        d = {name: m.trace_dump(W, H, mtx, eye, tm, max_bounces=bounces) for name, m in models.items()}
        a, b = d["rcp"]["hit_id"], d["ieee"]["hit_id"]
        reached = (a != -2) | (b != -2)
        for j in range(bounces):
            total[j] += int(reached[..., j].sum())
            same[j] += int(((a[..., j] == b[..., j]) & reached[..., j]).sum())
    res["hit_id_agreement_per_bounce"] = [float(s) / max(1, int(n)) for s, n in zip(same, total)]
    res["hit_id_mismatches_per_bounce"] = [int(n - s) for s, n in zip(same, total)]
    res["paths_compared"] = int(total[0])
    # ---- converged images
    acc = {name: np.zeros((H, W, 3), np.float32) for name in models}
    t0 = time.time()
    for k in range(spp):
        mtx, eye, w, tm = frame_uniforms(scene, k, W, H)
        for name, m in models.items():
            m.render_frame(W, H, mtx, eye, w, tm, acc[name], max_bounces=bounces, nthreads=threads)
    res["seconds"] = time.time() - t0
    a, b = acc["rcp"], acc["ieee"]
    diff = np.abs(a.astype(np.float64) - b.astype(np.float64))
    res["max_abs"] = float(diff.max())
    res["mean_abs"] = float(diff.mean())
    res["pixels_differing"] = int((diff.max(axis=2) > 0).sum())
    res["psnr_unclamped_db"] = psnr(a, b, max(1.0, float(a.max())))
    ac, bc = np.clip(a, 0, 1), np.clip(b, 0, 1)
    res["max_abs_clamped"] = float(np.abs(ac.astype(np.float64) - bc).max())
    res["psnr_clamped_db"] = psnr(ac, bc, 1.0)
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=512)
    ap.add_argument("--spp", type=int, default=1024)
    ap.add_argument("--configs", default="1,2,3")
    ap.add_argument("--threads", type=int, default=0)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    oracle_pkg.build()
    out = {"what": "a*rcp(b) model (oracle, CUDA) vs IEEE-division model, same frames", "results": []}
    for cfg in [int(c) for c in args.configs.split(",")]:
        r = compare(cfg, args.size, args.size, args.spp, threads=args.threads)
        print(json.dumps(r), flush=True)
        out["results"].append(r)
        if args.out:
            with open(args.out, "w") as f:
                json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
