"""Small renders of every kernel kind, to be run under compute-sanitizer (memcheck / racecheck / synccheck)."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
from sail_b200 import capi
from conftest import scene_tables, frame_uniforms

for cfg, kernel, name in ((1, 0, "mega"), (3, 1, "wavefront"), (2, 1, "wavefront"), (1, 3, "server"), (2, 3, "server")):
    W, H = 64, 40
    scene, o, t, l, bounces = scene_tables(cfg, W, H)
    with capi.Context(W, H) as ctx:
        ctx.configure(max_bounces=bounces, kernel=kernel)
        ctx.upload_scene(o, t, l)
        ctx.render_n([capi.make_frame(*frame_uniforms(scene, k, W, H)) for k in range(2)])
        img = ctx.read_accum()
        print(name, "cfg", cfg, ctx.kernel_info()[:60], float(img.mean()))
