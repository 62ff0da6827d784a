#!/usr/bin/env python
"""Both multi-GPU splits of SURVEY.md 8(e) on real GPUs, THROUGH THE C ABI (include/sail_b200.h: sail_create_info.ndev /
sail_group_join), checked against a single-GPU render of the same frames.  No torch, no torch.distributed.

    python tools/multigpu_check.py --gpus N [--out result.json]                 # one process, one context over N GPUs
    python tools/multigpu_check.py --gpus N --processes [--out result.json]     # N processes, one GPU each, joined group

  * sample sharding (config 4 style, Cornell scene): member g traces the frames k = g (mod N) into its own FP32 sum; the read /
    display calls combine the sums on the root (peer-memory slice sums, or ncclReduce).
    Bars: |sharded - single| <= 1e-5 * max(1, |single|) (summation order only); the peer-memory combine, whose order is
    fixed, equals the host-side sum of the members' buffers in member order bit for bit.
  * tile sharding (config 5: glass + metal, 16 bounces): member g traces interleaved 16-row bands; the rows are gathered on
    the root, then the display filter (box window, r = 2 px) runs there.
    Bar: accumulator, AOVs and displayed image bit-identical to the single-GPU ones.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from sail_b200 import capi, configs  # noqa: E402
from sail_b200.renderer import Renderer  # noqa: E402


def frames_of(r, scene, ks):
    return capi.frame_array([r.tracer.frame(scene.mat, scene.eye, k) for k in ks])


def single_gpu(scene_fn, W, H, NF, filt=None, device=0):
    scene, _, _, _, bounces = scene_fn(W, H)
    if filt:
        scene.filter = filt
        scene.filter.addParam("r", "vec2(2.0,2.0)")
    r = Renderer((W, H), device=device)
    r.update(scene)
    r.ctx.configure(max_bounces=bounces)
    r.ctx.render_n(frames_of(r, scene, range(NF)))
    out = {"accum": r.ctx.read_accum(), "hit": r.ctx.read_aov(capi.AOV_HIT_ID), "normal": r.ctx.read_aov(capi.AOV_NORMAL),
           "shown": r.ctx.present(as_u8=True).copy()}
    r.close()
    return out


def member_sums(scene_fn, W, H, NF, n):
    """What each member of an n-way sample-sharded context holds: frames k = g (mod n) in SUM mode, on one GPU."""
    scene, _, _, _, bounces = scene_fn(W, H)
    sums = []
    for g in range(n):
        r = Renderer((W, H), device=0)
        r.update(scene)
        r.ctx.configure(max_bounces=bounces, accum_mode=capi.ACCUM_SUM)
        r.ctx.render_n(frames_of(r, scene, [k for k in range(NF) if k % n == g]))
        sums.append(r.ctx.read_accum())
        r.close()
    return sums


def run_group(args, make_renderer, n, res, tag, is_root=True):
    """The checks, given a factory for the sharded renderer (in-process context or joined group member)."""
    W, H, NF = args.width, args.height, args.frames
    # ---------------- sample sharding ----------------
    scene, _, _, _, bounces = configs.config1(W, H)
    r = make_renderer(capi.SHARD_SAMPLE)
    r.update(scene)
    r.ctx.configure(max_bounces=bounces)
    fr = frames_of(r, scene, range(NF))
    r.ctx.render_n(capi.frame_array(list(fr)[:NF // 2]))   # two calls: the member assignment continues across calls
    r.ctx.render_n(capi.frame_array(list(fr)[NF // 2:]))
    got = r.ctx.read_accum()
    hit = r.ctx.read_aov(capi.AOV_HIT_ID)
    info = r.ctx.group_info()
    if is_root:
        want = single_gpu(configs.config1, W, H, NF)
        err = np.abs(got - want["accum"]) / np.maximum(1.0, np.abs(want["accum"]))
        res[tag + "sample_max_rel_err"] = float(err.max())
        res[tag + "sample_ok"] = bool(err.max() <= 1e-5) and bool(np.array_equal(hit, want["hit"]))
        res[tag + "info"] = info[2]
        res[tag + "combine_ms"] = r.ctx.last_combine_ms()
        if "p2p" in info[2]:
            sums = member_sums(configs.config1, W, H, NF, n)
            tot = sums[0].copy()
            for s in sums[1:]:
                tot = tot + s
            tot = tot * np.float32(1.0 / np.float32(NF))
            res[tag + "sample_p2p_bit_exact_vs_ordered_host_sum"] = bool(np.array_equal(tot.view(np.uint32), got.view(np.uint32)))
            res[tag + "sample_ok"] = res[tag + "sample_ok"] and res[tag + "sample_p2p_bit_exact_vs_ordered_host_sum"]
    r.close()
    # ---------------- tile sharding ----------------
    scene, _, _, _, bounces = configs.config5(W, H)
    scene.filter = "box"
    scene.filter.addParam("r", "vec2(2.0,2.0)")
    r = make_renderer(capi.SHARD_TILE)
    r.update(scene)
    r.ctx.configure(max_bounces=bounces)
    r.ctx.render_n(frames_of(r, scene, range(NF)))
    got = r.ctx.read_accum()
    nrm = r.ctx.read_aov(capi.AOV_NORMAL)
    shown = r.ctx.present(as_u8=True)
    if is_root:
        want = single_gpu(configs.config5, W, H, NF, filt="box")
        res[tag + "tile_accum_identical"] = bool(np.array_equal(got.view(np.uint32), want["accum"].view(np.uint32)))
        res[tag + "tile_aov_identical"] = bool(np.array_equal(nrm.view(np.uint32), want["normal"].view(np.uint32)))
        res[tag + "tile_display_identical"] = bool(np.array_equal(shown, want["shown"]))
        res[tag + "tile_ok"] = res[tag + "tile_accum_identical"] and res[tag + "tile_aov_identical"] and res[tag + "tile_display_identical"]
    r.close()


def rank_main(args):
    """One member of a joined group (spawned by --processes)."""
    rank, n = args.rank, args.gpus
    if rank == 0:
        gid = capi.group_id() + capi.group_id()  # one rendezvous token per group: sample-sharded, then tile-sharded
        with open(args.idfile + ".tmp", "wb") as f:
            f.write(gid)
        os.replace(args.idfile + ".tmp", args.idfile)
    else:
        t0 = time.time()
        while not os.path.exists(args.idfile):
            if time.time() - t0 > 120:
                raise RuntimeError("rank 0 never published the group id")
            time.sleep(0.05)
        gid = open(args.idfile, "rb").read()
    res = {"world": n, "mode": "processes"}

    def make(shard):
        token = gid[:128] if shard == capi.SHARD_SAMPLE else gid[128:]
        return Renderer((args.width, args.height), device=rank, group=(n, rank, token), shard=shard, band_rows=16)

    run_group(args, make, n, res, "", is_root=(rank == 0))
    if rank == 0:
        res["ok"] = bool(res["sample_ok"] and res["tile_ok"])
        with open(args.out, "w") as f:
            f.write(json.dumps(res) + "\n")
        print(json.dumps(res))
        return 0 if res["ok"] else 1
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=2)
    ap.add_argument("--processes", action="store_true")
    ap.add_argument("--out", default="")
    ap.add_argument("--width", type=int, default=640)
    ap.add_argument("--height", type=int, default=360)   # not a multiple of the 16-row band: the last band is short
    ap.add_argument("--frames", type=int, default=32)
    ap.add_argument("--rank", type=int, default=-1)
    ap.add_argument("--idfile", default="")
    args = ap.parse_args()
    if args.rank >= 0:
        sys.exit(rank_main(args))
    n = args.gpus
    if args.processes:
        tmp = tempfile.mkdtemp()
        idfile = os.path.join(tmp, "group.id")
        out = args.out or os.path.join(tmp, "out.json")
        procs = [subprocess.Popen([sys.executable, os.path.abspath(__file__), "--gpus", str(n), "--rank", str(k), "--idfile", idfile,
                                   "--out", out, "--width", str(args.width), "--height", str(args.height), "--frames", str(args.frames)])
                 for k in range(n)]
        rcs = [p.wait(timeout=540) for p in procs]
        if any(rcs):
            print("ranks exited with", rcs)
            sys.exit(1)
        print(open(out).read().strip())
        return
    res = {"world": n, "mode": "one process"}
    ok = True
    for combine, tag in ((capi.COMBINE_P2P, "p2p_"), (capi.COMBINE_NCCL, "nccl_")):
        def make(shard, combine=combine):
            return Renderer((args.width, args.height), devices=list(range(n)), shard=shard, combine=combine, band_rows=16)
        try:
            run_group(args, make, n, res, tag)
            ok = ok and res[tag + "sample_ok"] and res[tag + "tile_ok"]
        except capi.SailError as e:
            res[tag + "error"] = str(e)
            ok = False
    res["ok"] = bool(ok)
    line = json.dumps(res)
    print(line)
    if args.out:
        with open(args.out, "w") as f:
            f.write(line + "\n")
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
