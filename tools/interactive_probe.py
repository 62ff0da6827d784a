"""Where a 512x512 one-sample frame spends its time (bench.py's `interactive` leg, split): python tools/interactive_probe.py"""
import json
import sys
import time

sys.path.insert(0, ".")
import bench  # noqa: E402
from sail_b200 import capi, configs  # noqa: E402
from sail_b200.renderer import Renderer  # noqa: E402

out = {"interactive": bench.interactive_leg(0)}
scene, W, H, _, bounces = configs.config1()
r = Renderer((W, H))
r.ctx.configure(max_bounces=bounces)
r.update(scene)
r.render(scene)
r.ctx.specialise_wait()
N = 500
prepared = [r.tracer.frames(scene.mat, scene.eye, i, 1) for i in range(N)]
for fr in prepared[:20]:
    r.ctx.render_n(fr)
r.ctx.synchronize()


def timed(fn):
    t0 = time.perf_counter()
    fn()
    r.ctx.synchronize()
    return (time.perf_counter() - t0) / N * 1e6


def trace_async():
    for fr in prepared:
        r.ctx.render_n(fr)


def trace_sync():
    for fr in prepared:
        r.ctx.render_n(fr)
        r.ctx.synchronize()


def present_only():
    for _ in range(N):
        r.ctx.present(as_u8=True, pinned=True)


def host_frames():
    for i in range(N):
        r.tracer.frames(scene.mat, scene.eye, i, 1)


out["us_per_frame"] = {"trace_launch_async": timed(trace_async), "trace_launch_and_sync": timed(trace_sync),
                       "present_and_readback": timed(present_only), "python_frame_setup": timed(host_frames)}
print(json.dumps(out))
