"""Device-resident throughput against the number of frames per sail_render_n() launch (the launch's ramp-up and tail are a
fixed cost): python tools/batch_probe.py [workload ...]"""
import json
import sys

sys.path.insert(0, ".")
import bench  # noqa: E402

pk = bench.peaks()
import torch  # noqa: E402
peak = torch.cuda.get_device_properties(0).multi_processor_count * 128 * pk["sm_max_mhz"] * 1e6 / 1e12
out = {}
for wl in (sys.argv[1:] or ["cornell"]):
    for n in (8, 16, 32, 64, 128):
        bench.SPP_PER_STEP = n
        r = bench.other_workload(wl, 0, peak, steps=12 if n <= 32 else 6)
        out["%s/%d" % (wl, n)] = {"value": r["value"], "ms_per_step": r["ms_per_step"], "frac": r["roofline_frac"]}
        print(wl, n, "%.1f Mpaths/s  %.3f ms  frac %.3f" % (r["value"], r["ms_per_step"], r["roofline_frac"]), flush=True)
print(json.dumps(out))
