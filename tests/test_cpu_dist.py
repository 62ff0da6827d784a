"""world_size-2 `gloo` tests of the multi-GPU host logic (sail_b200/sharding.py) on CPU.

The per-rank renderer is the CPU oracle here (a test stand-in for the CUDA context): what is under test is
the work split and the combine step -- frame ownership k = rank (mod world), SUM-reduce + scale vs the
single-process running mean, and the row-band gather being a pure permutation."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT, frame_uniforms, scene_tables

W, H, NFRAMES = 40, 32, 6


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _render_sum(orc, scene, frames, rows=None, bounces=5):
    """Sum of the raw radiance samples of `frames` (what SAIL_ACCUM_SUM accumulates)."""
    total = np.zeros((H, W, 3), np.float32)
    for k in frames:
        m, eye, w, tm = frame_uniforms(scene, k, W, H)
        total += orc.trace_dump(W, H, m, eye, tm, max_bounces=bounces)["radiance"]
    return total


def _worker(rank, world, port, out_path):
    import sys
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle as oracle_pkg
    from sail_b200 import sharding
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        scene, o, t, l, bounces = scene_tables(2, W, H)
        orc = oracle_pkg.Oracle()
        orc.set_scene(o, t, l)
        # ---- sample sharding ----
        mine = sharding.frames_for_rank(0, NFRAMES, rank, world)
        part = torch.from_numpy(_render_sum(orc, scene, mine, bounces=bounces))
        sharding.reduce_sum(part, dst=0)
        # ---- tile sharding: every rank renders all frames for its bands only ----
        bands = sharding.row_bands(H, world, band=8)
        full = _render_sum(orc, scene, range(NFRAMES), bounces=bounces)
        masked = np.full_like(full, np.nan)
        for a, b in bands[rank]:
            masked[a:b] = full[a:b]
        gathered = sharding.gather_rows(torch.from_numpy(masked), bands[rank], bands, dst=0)
        if rank == 0:
            np.savez(out_path, summed=sharding.mean_from_sum(part.numpy(), NFRAMES), gathered=gathered.numpy(), full=full)
    finally:
        dist.destroy_process_group()


def test_frame_and_row_ownership():
    from sail_b200 import sharding
    for world in (1, 2, 4, 8):
        owned = [sharding.frames_for_rank(3, 29, r, world) for r in range(world)]
        assert sorted(sum(owned, [])) == list(range(3, 32))
        bands = sharding.row_bands(1080, world, band=16)
        rows = sorted(y for r in bands for a, b in r for y in range(a, b))
        assert rows == list(range(1080))
        stripes = [sharding.contiguous_rows(1080, r, world) for r in range(world)]
        assert stripes[0][0] == 0 and stripes[-1][1] == 1080 and all(a % 8 == 0 for a, _ in stripes)
        assert all(stripes[i][1] == stripes[i + 1][0] for i in range(world - 1))


@pytest.mark.timeout(300)
def test_two_rank_gloo_sample_and_tile_sharding(oracle, tmp_path):
    out = str(tmp_path / "rank0.npz")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    r = np.load(out)
    # single-process reference: the running mean of fstrace.glsl:14 over the same frames
    scene, o, t, l, bounces = scene_tables(2, W, H)
    oracle.set_scene(o, t, l)
    acc = np.zeros((H, W, 3), np.float32)
    for k in range(NFRAMES):
        m, eye, w, tm = frame_uniforms(scene, k, W, H)
        oracle.render_frame(W, H, m, eye, w, tm, acc, max_bounces=bounces)
    # summation order differs from the running mean: 1e-5 relative-or-absolute
    assert np.all(np.abs(r["summed"] - acc) <= 1e-5 * np.maximum(1.0, np.abs(acc)))
    # tile sharding moves rows, no arithmetic: bit-identical
    assert np.array_equal(r["gathered"].view(np.uint32), r["full"].view(np.uint32))
