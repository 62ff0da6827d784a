#!/usr/bin/env python
"""The one exchange step of the path, timed alone: combining the members' accumulation buffers of a sharded context on its
root, peer-memory kernels / copies (sail_combine.cuh) against ncclReduce, one process over N GPUs.

    python tools/combine_bench.py --gpus N [--out profiles/rNN_combine.json]

Per size (1080p, 4K) and shard mode: device time on the root from the end of its own trace to the combined image
(sail_last_combine_ms), median of `--iters` warmed calls; GB/s = image bytes (W*H*12) / that time.
"""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from sail_b200 import capi, configs  # noqa: E402
from sail_b200.renderer import Renderer  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=2)
    ap.add_argument("--iters", type=int, default=12)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    n = args.gpus
    res = {"gpus": n, "runs": []}
    for (W, H) in ((1920, 1080), (3840, 2160)):
        for shard, sname in ((capi.SHARD_SAMPLE, "sample"), (capi.SHARD_TILE, "tile")):
            for combine, cname in ((capi.COMBINE_P2P, "p2p"), (capi.COMBINE_NCCL, "nccl")):
                scene, _, _, _, bounces = configs.config1(W, H)
                try:
                    r = Renderer((W, H), devices=list(range(n)), shard=shard, combine=combine)
                except capi.SailError as e:
                    res["runs"].append({"size": [W, H], "shard": sname, "combine": cname, "error": str(e)})
                    continue
                r.update(scene)
                r.ctx.configure(max_bounces=bounces)
                ms = []
                for i in range(args.iters + 3):
                    r.trace(scene, frames=n)            # one frame per member, then ask for the image
                    r.ctx.present(as_u8=True, pinned=True)
                    if i >= 3:
                        ms.append(r.ctx.last_combine_ms())
                info = r.ctx.group_info()[2]
                r.close()
                med = float(np.median(ms))
                res["runs"].append({"size": [W, H], "shard": sname, "combine": cname, "info": info, "ms_median": med, "ms_min": float(min(ms)),
                                    "image_mb": W * H * 12 / 1e6, "gbs": W * H * 12 / (med * 1e-3) / 1e9})
                print(res["runs"][-1], flush=True)
    if args.out:
        with open(args.out, "w") as f:
            json.dump(res, f, indent=1)


if __name__ == "__main__":
    main()
