#!/usr/bin/env python
"""Wall time of the display pass (filter kernel + D2H of the RGBA8 image into page-locked memory) per filter at 1080p / 4K."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from sail_b200 import capi, configs  # noqa: E402
from sail_b200.filters import lower_filter  # noqa: E402
from sail_b200.generator import PluginParams  # noqa: E402
from sail_b200.renderer import Renderer  # noqa: E402

out = []
for (W, H) in ((1920, 1080), (3840, 2160)):
    scene, _, _, _, bounces = configs.config5(W, H)
    r = Renderer((W, H))
    r.update(scene)
    r.trace(scene, frames=4)
    for name in ("color", "tonemapping", "box", "sinc", "gaussian", "wavelet"):
        pp = PluginParams(name)
        for k, v in (("r", "vec2(2.0,2.0)"), ("tau", "3.0"), ("alpha", "2.0")):
            pp.addParam(k, v)
        f = lower_filter(pp)
        r.ctx.configure(filter_id=f["filter_id"], filter_gamma_c=f["gamma_c"], filter_radius=f["radius"], filter_weights=f["weights"])
        for _ in range(3):
            r.ctx.present(as_u8=True, pinned=True)
        t0 = time.perf_counter()
        for _ in range(20):
            r.ctx.present(as_u8=True, pinned=True)
        ms = (time.perf_counter() - t0) / 20 * 1e3
        out.append({"size": [W, H], "filter": name, "present_ms": ms})
        print(out[-1], flush=True)
    r.close()
if len(sys.argv) > 1:
    json.dump(out, open(sys.argv[1], "w"), indent=1)
