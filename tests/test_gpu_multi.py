"""Multi-GPU sharding on real devices, driven through the C ABI: sail_create_info.ndev (one context over
several GPUs of this process; peer-memory and NCCL combine) and sail_group_join (one GPU per process).  Skipped on a one-GPU
box.  The host-side split logic has a CPU twin in tests/test_cpu_dist.py."""
import json
import os
import subprocess
import sys

import pytest

from conftest import ROOT


def _gpus():
    from sail_b200 import capi
    return min(capi.device_count(), 4)


@pytest.mark.gpu
@pytest.mark.timeout(600)
@pytest.mark.parametrize("mode", ["one_process", "processes"])
def test_sample_and_tile_sharding_through_the_c_abi(tmp_path, mode):
    n = _gpus()
    if n < 2:
        pytest.skip("needs >= 2 GPUs (found %d)" % n)
    out = str(tmp_path / "multi.json")
    cmd = [sys.executable, os.path.join(ROOT, "tools", "multigpu_check.py"), "--gpus", str(n), "--out", out]
    if mode == "processes":
        cmd.append("--processes")
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=570)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-4000:]
    res = json.load(open(out))
    assert res["world"] == n and res["ok"], res


@pytest.mark.gpu
def test_sharded_context_argument_checks():
    from sail_b200 import capi
    with pytest.raises(capi.SailError) as ei:  # two members may share a GPU, but NCCL cannot put two ranks on one device
        capi.Context(64, 64, devices=[0, 0], shard=capi.SHARD_SAMPLE, combine=capi.COMBINE_NCCL)
    assert ei.value.code == capi.ERR_INVALID and "distinct devices" in str(ei.value)
    with pytest.raises(capi.SailError):
        capi.Context(64, 64, devices=[0, 99], shard=capi.SHARD_TILE)
    with capi.Context(64, 64) as ctx:
        assert ctx.group_info()[:2] == (1, 0)
        with pytest.raises(capi.SailError):
            ctx.last_combine_ms()
        with pytest.raises(capi.SailError):
            ctx.group_join(1, 0, bytes(128))


@pytest.mark.gpu
@pytest.mark.parametrize("members", [2, 3, 8])
def test_sharded_context_with_members_sharing_one_gpu(members):
    """The whole sharded path -- dealing frames / row bands to members, per-member streams and buffers, the peer-memory combine
    kernels, AOVs on the root, the combined display image -- on ONE GPU: `device_ids` may list a device several times.  Same
    checks as on real multi-GPU boxes (tools/multigpu_check.py): sample-sharded within summation order of a single context and
    bit-identical to the ordered host sum of the members' buffers; tile-sharded accumulator, AOVs and display bit-identical."""
    import argparse
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import multigpu_check as M
    from sail_b200 import capi
    from sail_b200.renderer import Renderer
    args = argparse.Namespace(width=320, height=200, frames=16)  # 200 rows: the last 16-row band is short
    res = {}
    M.run_group(args, lambda shard: Renderer((args.width, args.height), devices=[0] * members, shard=shard, combine=capi.COMBINE_P2P,
                                             band_rows=16), members, res, "")
    assert res["sample_ok"] and res["sample_p2p_bit_exact_vs_ordered_host_sum"] and res["tile_ok"], res
    assert "p2p" in res["info"] and str(members) in res["info"]


@pytest.mark.gpu
def test_sample_sharded_restart_reset_and_batches_follow_the_single_context():
    """A sample-sharded context behaves like one context through the calls a host makes: a frame with textureWeight = 0 restarts
    the running mean (renderer.js:57-60), sail_reset clears it, a batch larger than 32 frames is dealt out over the members."""
    import numpy as np
    from conftest import frame_uniforms, scene_tables
    from sail_b200 import capi
    W, H = 96, 64
    scene, o, t, l, bounces = scene_tables(2, W, H)

    def frames(ks):
        return [capi.make_frame(*frame_uniforms(scene, k, W, H)) for k in ks]

    def drive(ctx):
        out = []
        ctx.configure(max_bounces=bounces)
        ctx.upload_scene(o, t, l)
        ctx.render_n(frames(range(0, 5)))
        out.append(ctx.read_accum())
        ctx.render_n(frames([5, 6, 0, 1, 2]))       # restart in the middle of a batch: frames 5, 6 are discarded by the mean
        out.append(ctx.read_accum())
        ctx.render_n(frames([3]))
        out.append(ctx.read_accum())
        out.append(ctx.read_accum())                 # cached combine: same image
        ctx.reset()
        ctx.render_n(frames(range(0, 70)))           # more than SAIL_MAX_FRAMES_PER_CALL: chunked by the binding
        out.append(ctx.read_accum())
        out.append(ctx.read_aov(capi.AOV_HIT_ID).astype(np.float32))
        return out

    with capi.Context(W, H) as single:
        want = drive(single)
    with capi.Context(W, H, devices=[0, 0, 0], shard=capi.SHARD_SAMPLE) as sharded:
        got = drive(sharded)
    for i, (g, w) in enumerate(zip(got, want)):
        assert np.all(np.abs(g - w) <= 1e-5 * np.maximum(1.0, np.abs(w))), "step %d: max abs diff %g" % (i, np.abs(g - w).max())
    assert np.array_equal(got[2], got[3]) and np.array_equal(got[5], want[5])
