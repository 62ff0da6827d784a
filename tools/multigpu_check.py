#!/usr/bin/env python
"""Both multi-GPU splits of SURVEY.md 8(e) on real GPUs, checked against a single-GPU render of the same frames.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        tools/multigpu_check.py [--out result.json]

One process per GPU, NCCL over NVLink.  Rank 0 additionally renders everything alone (the reference result).
  * sample sharding (config 4 style, Cornell scene): rank g traces frames k = g (mod N) in SAIL_ACCUM_SUM mode, one NCCL
    reduce(SUM) of the accumulator (wrapped zero-copy through __cuda_array_interface__), scale by 1/n.
    Bar: |sharded - single| <= 1e-5 * max(1, |single|)  (summation order only).
  * tile sharding (config 5: glass + metal, 16 bounces): rank g traces interleaved 16-row bands (sail_set_row_bands),
    one NCCL gather of the rows, then the display filter (box window, r = 2 px) on the root.
    Bar: accumulator and displayed image bit-identical to the single-GPU ones.
"""
import argparse
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from sail_b200 import capi, configs, sharding  # noqa: E402
from sail_b200.renderer import Renderer  # noqa: E402


def frames_of(r, scene, ks):
    return capi.frame_array([r.tracer.frame(scene.mat, scene.eye, k) for k in ks])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="")
    ap.add_argument("--width", type=int, default=640)
    ap.add_argument("--height", type=int, default=368)
    ap.add_argument("--frames", type=int, default=32)
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    W, H, NF = args.width, args.height, args.frames
    res = {"world": world, "width": W, "height": H, "frames": NF}

    # ---------------- sample sharding ----------------
    scene, _, _, _, bounces = configs.config1(W, H)
    r = Renderer((W, H), device=local)
    r.ctx.configure(max_bounces=bounces, accum_mode=capi.ACCUM_SUM)
    r.update(scene)
    r.ctx.reset()
    r.ctx.render_n(frames_of(r, scene, sharding.frames_for_rank(0, NF, rank, world)))
    r.ctx.synchronize()
    acc = torch.as_tensor(sharding.DeviceBuffer(r.ctx.device_ptr(capi.AOV_COLOR), (H, W, 3)), device="cuda")
    sharding.reduce_sum(acc, dst=0)
    torch.cuda.synchronize()
    if rank == 0:
        r.ctx.scale_accum(1.0 / NF)
        got = r.ctx.read_accum()
        r.ctx.configure(accum_mode=capi.ACCUM_MEAN)
        r.ctx.reset()
        r.ctx.render_n(frames_of(r, scene, range(NF)))
        want = r.ctx.read_accum()
        err = np.abs(got - want) / np.maximum(1.0, np.abs(want))
        res["sample_max_rel_err"] = float(err.max())
        res["sample_ok"] = bool(err.max() <= 1e-5)
    r.close()

    # ---------------- tile sharding ----------------
    scene, _, _, _, bounces = configs.config5(W, H)
    scene.filter = "box"
    scene.filter.addParam("r", "vec2(2.0,2.0)")
    r = Renderer((W, H), device=local)
    r.update(scene)
    r.ctx.configure(max_bounces=bounces)
    bands = sharding.row_bands(H, world, band=16)
    r.ctx.set_row_bands(16, world, rank)
    r.ctx.reset()
    fr = frames_of(r, scene, range(NF))
    r.ctx.render_n(fr)
    r.ctx.synchronize()
    acc = torch.as_tensor(sharding.DeviceBuffer(r.ctx.device_ptr(capi.AOV_COLOR), (H, W, 3)), device="cuda")
    full = sharding.gather_rows(acc, bands[rank], bands, dst=0)
    torch.cuda.synchronize()
    if rank == 0:
        gathered = full.cpu().numpy()
        r.ctx.set_row_bands(16, 1, 0)               # root takes the whole image back
        r.ctx.write_accum(gathered)
        shown = r.ctx.present(as_u8=True).copy()    # display filter after the gather (needs neighbours across bands)
        r.ctx.reset()
        r.ctx.render_n(fr)
        want = r.ctx.read_accum()
        want8 = r.ctx.present(as_u8=True)
        res["tile_accum_identical"] = bool(np.array_equal(gathered.view(np.uint32), want.view(np.uint32)))
        # AOVs (normal / position feed only the wavelet filter) are not gathered: the box filter reads colour only
        res["tile_display_identical"] = bool(np.array_equal(shown, want8))
    r.close()
    dist.barrier()
    if rank == 0:
        res["ok"] = bool(res["sample_ok"] and res["tile_accum_identical"] and res["tile_display_identical"])
        line = json.dumps(res)
        print(line)
        if args.out:
            with open(args.out, "w") as f:
                f.write(line + "\n")
    dist.destroy_process_group()
    if rank == 0 and not res["ok"]:
        sys.exit(1)


if __name__ == "__main__":
    main()
