#!/usr/bin/env python
"""A small tour of every kernel for compute-sanitizer (memcheck / racecheck / initcheck):
    compute-sanitizer --tool racecheck python tools/sanitize_run.py
Both trace kernels (scene-specialised and ahead-of-time) on configs 1, 3 and 5, the on-demand AOV kernel, the three display kernels,
the diagnostics, the fixed-scratch intersection query and a sample- and a tile-sharded context whose members share the GPU
(peer-memory combine kernels / band copies)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
os.environ.setdefault("SAIL_B200_JIT_SYNC", "1")
from conftest import frame_uniforms, scene_tables  # noqa: E402
from sail_b200 import capi  # noqa: E402
from sail_b200.filters import lower_filter  # noqa: E402
from sail_b200.generator import PluginParams  # noqa: E402

W, H = 64, 48
launches = 0
for cfg in (1, 3, 5):
    scene, o, t, l, bounces = scene_tables(cfg, W, H)
    frames = [capi.make_frame(*frame_uniforms(scene, k, W, H)) for k in range(2)]
    for kernel in (capi.KERNEL_MEGA, capi.KERNEL_WAVEFRONT):
        for spec in (capi.SPECIALISE_REQUIRE, capi.SPECIALISE_OFF):
            with capi.Context(W, H) as ctx:
                ctx.configure(max_bounces=min(bounces, 6), kernel=kernel, specialise=spec, rr_start_bounce=3 if cfg == 5 else 0)
                ctx.upload_scene(o, t, l)
                ctx.render_n(frames)
                ctx.read_accum()
                ctx.read_aov(capi.AOV_NORMAL)
                for name in ("tonemapping", "gaussian", "wavelet"):
                    pp = PluginParams(name)
                    pp.addParam("r", "vec2(2.0,2.0)")
                    pp.addParam("alpha", "2.0")
                    f = lower_filter(pp)
                    ctx.configure(filter_id=f["filter_id"], filter_gamma_c=f["gamma_c"], filter_radius=f["radius"], filter_weights=f["weights"])
                    ctx.present(both=True)
                rays = np.random.default_rng(0).normal(size=(500, 6)).astype(np.float32)
                ctx.intersect_rays(rays)
                ctx.intersect_rays_noalloc(rays)
                ctx.trace_dump(frames[0])
                launches += ctx.launch_count()
    for shard in (capi.SHARD_SAMPLE, capi.SHARD_TILE):
        with capi.Context(W, H, devices=[0, 0, 0], shard=shard, band_rows=8) as ctx:
            ctx.configure(max_bounces=min(bounces, 6))
            ctx.upload_scene(o, t, l)
            ctx.render_n(frames + frames)
            ctx.present(as_u8=True)
            ctx.read_aov(capi.AOV_HIT_ID)
            launches += ctx.launch_count()
print("sanitize_run: done, %d kernel launches" % launches)
