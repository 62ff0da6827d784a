#!/usr/bin/env python
"""bench.py -- headline benchmark: Mpaths/s on the 1080p Cornell box (BASELINE.json `metric`).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" = SPP_PER_STEP progressive samples per pixel of the 1920x1080 README Cornell box (5 bounces) on
every rank = one sail_render_n() launch per rank.  Multi-GPU is sample sharding: rank g
traces frame indices k = g (mod N) into a local FP32 sum; nothing is exchanged per step, so scaling is
"weak" (per-GPU work fixed).  After the timed region the partial sums are combined once with an NCCL
reduce (its time is reported as `reduce_ms`) and the result sanity-checked on rank 0.

value  : paths/s with everything resident in HBM, CUDA events on the context's stream, max over ranks.
e2e    : same metric through sail_b200.Renderer (the user-facing call): every step builds the frame
         uniforms on the host, launches, runs the display filter and copies the RGBA8 image to the host.
         `e2e.value` uses Renderer.render_pipelined (display double-buffered: the copy of step i overlaps the trace
         of step i+1, all copies inside the timed region), `e2e.synchronous_value` uses Renderer.render.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOAD = "README Cornell box 1920x1080, 5 bounces (config 1 scene at 1080p)"
SPP_PER_STEP = 32
METRIC, UNIT = "Mpaths/s", "Mpaths/s"


def peaks():
    p = {"sm_max_mhz": 1965.0, "hbm_gbs": 6650.0, "source": "fallback"}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            m = json.load(f)
        p.update(sm_max_mhz=float(m["sm_max_mhz"]), hbm_gbs=float(m["hbm_gbs"]), source="measured")
    except Exception:
        pass
    return p


def ops_per_path():
    key = {"cornell": "cornell_1080p", "config2": "config2_1080p", "config3": "config3_1080p", "config5": "config5_1080p",
           "config4": "config4_2160p"}[_workload]
    with open(os.path.join(ROOT, "baseline", "ops_per_path.json")) as f:
        return float(json.load(f)[key]["ops_per_path"])


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
            "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = max(mx, float(r[1]))
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


WORKLOADS = {"cornell": ("cornell_1080p", WORKLOAD),
             "config2": ("config2", "materials demo (mirror/glass/metal/matte spheres, area light) 1920x1080, 5 bounces"),
             "config3": ("config3", "quadrics + textures + area/point/spot lights 1920x1080, 5 bounces"),
             "config5": ("config5", "glass + metal spheres 1920x1080, 16 bounces"),
             "config4": ("config4", "Cornell box 3840x2160, 5 bounces, sample-sharded across the GPUs (BASELINE config 4)")}
_workload = "cornell"


def build_workload():
    from sail_b200 import configs, wire
    scene, W, H, _, bounces = getattr(configs, WORKLOADS[_workload][0])()
    o, t, l = wire.serialize_scene(scene)
    return scene, W, H, bounces, (o, t, l)


def other_workload(name, local_rank, peak_tops, steps=8, warm=3):
    """Device-resident throughput of one of the other BASELINE.json configs (same timing rules, fewer steps):
    context for the headline, not a bench line of its own."""
    global _workload
    import torch
    from sail_b200 import capi
    from sail_b200.renderer import Renderer
    keep, _workload = _workload, name
    try:
        scene, W, H, bounces, _ = build_workload()
        ops = ops_per_path()
    finally:
        _workload = keep
    r = Renderer((W, H), device=local_rank)
    r.ctx.configure(max_bounces=bounces)
    r.update(scene)
    batches = [r.tracer.frames(scene.mat, scene.eye, s * SPP_PER_STEP, SPP_PER_STEP) for s in range(warm + steps)]
    stream = torch.cuda.ExternalStream(r.ctx.stream(), device=local_rank)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for s in range(warm):
        r.ctx.render_n(batches[s])
    r.ctx.synchronize()
    with torch.cuda.stream(stream):
        for s in range(steps):
            flush.zero_()
            ev[s][0].record(stream)
            r.ctx.render_n(batches[warm + s])
            ev[s][1].record(stream)
    r.ctx.synchronize()
    torch.cuda.synchronize()
    ms = sum(a.elapsed_time(b) for a, b in ev) / steps
    feats = capi.scene_features(*_tables(scene))
    used = r.ctx.kernel_info()
    r.close()
    paths_s = W * H * SPP_PER_STEP / (ms * 1e-3)
    return {"workload": WORKLOADS[name][1], "value": paths_s / 1e6, "unit": UNIT, "ms_per_step": ms, "steps": steps,
            "kernel_variant_F": int(feats), "kernel_used": used, "roofline_frac": paths_s * ops / 1e12 / peak_tops, "ops_per_path": ops}


def _tables(scene):
    from sail_b200 import wire
    return wire.serialize_scene(scene)


def measured_traffic():
    """DRAM bytes per launch of the dominant kernel from the committed ncu capture (profiles/traffic.json), or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f).get(_workload, {}).get("dram_bytes_per_launch")
    except (OSError, ValueError):
        return None


def measured_ncu():
    """Pipe-side counters of the same ncu capture (issue-slot utilisation, executed instructions), or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f).get(_workload, {}).get("ncu")
    except (OSError, ValueError):
        return None


def cpu_baseline(seconds=12.0, min_frames=8, threads=0):
    """The CPU oracle (a port: the reference has no CPU implementation) on a bounded sample of the workload: whole
    1080p frames of 1 spp until about ``seconds`` of CPU work are done, all host cores (OpenMP over rows)."""
    import oracle as oracle_pkg
    from sail_b200 import schedule
    scene, W, H, bounces, (o, t, l) = build_workload()
    orc = oracle_pkg.Oracle()
    orc.set_scene(o, t, l)
    sch = schedule.FrameSchedule()
    acc = np.zeros((H, W, 3), np.float32)
    cores = threads or (os.cpu_count() or 1)
    total, frames, k = 0.0, 0, 0
    while frames < min_frames or total < seconds:
        m = schedule.frame_matrix(scene.mat, sch.jitter_unit(k), W, H)
        t0 = time.perf_counter()
        orc.render_frame(W, H, m, scene.eye, k / (k + 1.0), sch.time(k), acc, max_bounces=bounces, nthreads=cores)
        dt = time.perf_counter() - t0
        if k > 0:  # first frame = warm-up
            total += dt
            frames += 1
        k += 1
        if frames >= 400:
            break
    return {"value": frames * W * H / total / 1e6, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": "%d full 1080p frames of 1 spp (%d paths) in %.1f s, OpenMP over rows" % (frames, frames * W * H, total)}


def run_reference(args, rank, world):
    """Reference arm: the reference is GLSL/JS with no runnable CPU or GPU path on this box (no graphics
    stack, no Node.js); its timed stand-in is the CPU oracle -- the literal restatement of the shader math --
    on all host cores.  Rank 0 only."""
    if rank != 0:
        return
    steps, warm = args.steps, args.warmup
    import oracle as oracle_pkg
    from sail_b200 import schedule
    scene, W, H, bounces, (o, t, l) = build_workload()
    orc = oracle_pkg.Oracle()
    orc.set_scene(o, t, l)
    sch = schedule.FrameSchedule()
    acc = np.zeros((H, W, 3), np.float32)
    cores = os.cpu_count() or 1
    times = []
    for k in range(warm + steps):  # one step = one full 1080p frame (1 spp): the bounded sample of the workload
        m = schedule.frame_matrix(scene.mat, sch.jitter_unit(k), W, H)
        t0 = time.perf_counter()
        orc.render_frame(W, H, m, scene.eye, k / (k + 1.0), sch.time(k), acc, max_bounces=bounces, nthreads=cores)
        if k >= warm:
            times.append(time.perf_counter() - t0)
    total = sum(times)
    value = steps * W * H / total / 1e6
    sample = "each step = 1 spp over the full 1920x1080 frame (%d paths) on %d host threads" % (W * H, cores)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": steps, "warmup": warm,
        "ms_per_step": 1e3 * total / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "config": {"workload": WORKLOAD, "spp_per_step": 1, "note": "CPU oracle port; the reference's "
                                        "WebGL path is unavailable on this box"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}))


class _DevBuf:
    """__cuda_array_interface__ view of a device buffer owned by a sail_ctx (for torch.distributed)."""

    def __init__(self, ptr, shape):
        self.__cuda_array_interface__ = {"shape": shape, "typestr": "<f4", "data": (int(ptr), False), "version": 2}


def run_ours(args, rank, world, local_rank):
    import torch

    from sail_b200 import capi
    from sail_b200.renderer import Renderer
    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    if not torch.cuda.is_available() or capi.device_count() < 1:
        raise SystemExit("bench.py: no sm_100 GPU visible; the CUDA path has no fallback")
    torch.cuda.set_device(local_rank)
    scene, W, H, bounces, (o, t, l) = build_workload()
    steps, warm = args.steps, args.warmup
    paths_per_step = W * H * SPP_PER_STEP

    # ---------------- device-resident arm -----------------
    r = Renderer((W, H), device=local_rank)
    ctx = r.ctx
    ctx.configure(max_bounces=bounces, accum_mode=capi.ACCUM_SUM if world > 1 else capi.ACCUM_MEAN,
                  kernel={"wavefront": capi.KERNEL_WAVEFRONT, "mega": capi.KERNEL_MEGA, "auto": capi.KERNEL_AUTO, "server": capi.KERNEL_SERVER}[args.kernel],
                  specialise=capi.SPECIALISE_OFF if args.no_jit else capi.SPECIALISE_AUTO)
    r.update(scene)
    total_steps = warm + steps
    # rank g owns frame indices k = g (mod world)
    batches = []
    for s in range(total_steps):
        ks = [(s * SPP_PER_STEP + i) * world + rank for i in range(SPP_PER_STEP)]
        batches.append(capi.frame_array([r.tracer.frame(scene.mat, scene.eye, k) for k in ks]))
    stream = torch.cuda.ExternalStream(ctx.stream(), device=local_rank)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    sampler = ClockSampler(local_rank)
    sampler.start()  # runs through the warm-up (same kernel, same load) and the timed region
    for s in range(warm):
        ctx.render_n(batches[s])
    ctx.synchronize()
    torch.cuda.synchronize()
    if dist:
        dist.barrier()
    launches0 = ctx.launch_count()
    with torch.cuda.stream(stream):
        for s in range(steps):
            flush.zero_()  # L2 flush (256 MiB > 126 MB L2), outside the per-step events
            ev[s][0].record(stream)
            ctx.render_n(batches[warm + s])
            ev[s][1].record(stream)
    ctx.synchronize()
    torch.cuda.synchronize()
    launches = ctx.launch_count() - launches0
    kernel_used = ctx.kernel_info()
    clocks = sampler.stop()
    if dist:
        dist.barrier()
    dev_ms = sum(a.elapsed_time(b) for a, b in ev)

    # ---------------- multi-GPU combine (once per presentation, outside the per-step timing) -----------------
    reduce_ms = None
    if dist:
        acc = torch.as_tensor(_DevBuf(ctx.device_ptr(capi.AOV_COLOR), (H, W, 3)), device="cuda")
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        dist.reduce(acc, dst=0, op=dist.ReduceOp.SUM)
        e1.record()
        torch.cuda.synchronize()
        reduce_ms = e0.elapsed_time(e1)
        if rank == 0:
            ctx.scale_accum(1.0 / (total_steps * SPP_PER_STEP * world))
            img = ctx.read_accum()
            assert np.isfinite(img).all() and 0.0 < float(img.mean()) < 1.0, "reduced image is not sane"

    # ---------------- end-to-end arm: Renderer API, host buffers in and out -----------------
    scene.sampleCount = 0
    e2e_steps = max(1, min(steps, 20))
    for _ in range(2):
        r.render(scene, frames=SPP_PER_STEP)
    if dist:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        # host: 32 x (jitter, inverse MVP) -> sail_frame[] -> H2D as kernel parameters; trace; display filter;
        # D2H of the RGBA8 image into page-locked host memory (synchronises)
        img8 = r.render(scene, frames=SPP_PER_STEP)
    e2e_s = time.perf_counter() - t0
    assert img8.shape == (H, W, 4)
    # the same steps with the display double-buffered (Renderer.render_pipelined: step i's image is collected by call i+1,
    # its device-to-host copy overlaps the tracing of step i+1; every copy still lies inside the timed region)
    for _ in range(2):
        r.render_pipelined(scene, frames=SPP_PER_STEP)
    r.flush()
    if dist:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        img8 = r.render_pipelined(scene, frames=SPP_PER_STEP)
    img8 = r.flush()
    e2e_pipe_s = time.perf_counter() - t0
    assert img8.shape == (H, W, 4)

    times = torch.tensor([dev_ms, e2e_s * 1e3, e2e_pipe_s * 1e3], dtype=torch.float64, device="cuda")
    if dist:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms, e2e_pipe_ms = float(times[0]), float(times[1]), float(times[2])
    if rank == 0:
        pk = peaks()
        n_sm = torch.cuda.get_device_properties(local_rank).multi_processor_count
        value = world * steps * paths_per_step / (dev_ms * 1e-3) / 1e6
        per_gpu_paths_s = steps * paths_per_step / (dev_ms * 1e-3)
        ops = ops_per_path()
        peak = n_sm * 128 * pk["sm_max_mhz"] * 1e6 / 1e12
        achieved = per_gpu_paths_s * ops / 1e12
        hbm_gbs = (W * H * 24) / (dev_ms / steps * 1e-3) / 1e9
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": warm,
            "ms_per_step": dev_ms / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": WORKLOADS[_workload][1], "spp_per_step": SPP_PER_STEP, "paths_per_step_per_gpu": paths_per_step,
                       "kernel": args.kernel, "kernel_used": kernel_used, "arithmetic": "strict (bit-exact vs oracle)", "sharding": "sample index mod N",
                       "l2": "flushed between steps (256 MiB memset outside the per-step events)"},
            "spp_per_s": value * 1e6 / (W * H),
            "roofline": {"bound": "fp32_issue", "achieved": achieved, "peak": peak, "unit": "Tops/s (FP32 lane-ops)",
                         "frac": achieved / peak, "traffic": measured_traffic(), "ncu": measured_ncu(), "ops_per_path": ops,
                         "peak_def": "%d SMs x 128 lanes x %.0f MHz (%s sm_max_mhz)" % (n_sm, pk["sm_max_mhz"], pk["source"]),
                         "hbm": {"achieved": hbm_gbs, "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": hbm_gbs / pk["hbm_gbs"],
                                 "bytes_per_launch": W * H * 24}},
            "clocks": clocks,
            "e2e": {"value": world * e2e_steps * paths_per_step / (e2e_pipe_ms * 1e-3) / 1e6, "unit": UNIT,
                    "h2d_bytes_per_step": SPP_PER_STEP * 84, "d2h_bytes_per_step": W * H * 4, "steps": e2e_steps,
                    "mode": "Renderer.render_pipelined per step: host frame construction, H2D of the frame uniforms, trace, display "
                            "filter, D2H of the RGBA8 image into page-locked memory; the display is double-buffered (step i's image "
                            "is collected by call i+1, its copy overlaps the next trace), every copy lies inside the timed region",
                    "synchronous_value": world * e2e_steps * paths_per_step / (e2e_ms * 1e-3) / 1e6,
                    "synchronous_mode": "Renderer.render per step: the same work, the call returns its own image (waits for the copy)"},
            "gpu_launches": int(launches),
        }
        if reduce_ms is not None:
            line["reduce_ms"] = reduce_ms
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_baseline()
        if world == 1 and not args.no_cpu and _workload == "cornell":
            line["other_workloads"] = [other_workload(w, local_rank, peak) for w in ("config2", "config3", "config5")]
            # The same workload with the last-bounce reductions compiled out (every bounce evaluated in full, as the
            # reference's loop does; identical image): how much of `value` is skipped dead work.
            keep = os.environ.get("SAIL_B200_JIT_DEFINES")
            os.environ["SAIL_B200_JIT_DEFINES"] = (keep + "," if keep else "") + "SAIL_LITERAL_LAST_BOUNCE"
            try:
                lit = other_workload("cornell", local_rank, peak)
            finally:
                if keep is None:
                    del os.environ["SAIL_B200_JIT_DEFINES"]
                else:
                    os.environ["SAIL_B200_JIT_DEFINES"] = keep
            line["literal_last_bounce"] = {"value": lit["value"], "unit": UNIT, "ms_per_step": lit["ms_per_step"], "steps": lit["steps"],
                                           "roofline_frac": lit["roofline_frac"], "kernel_used": lit["kernel_used"],
                                           "note": "same workload, kernel built with -DSAIL_LITERAL_LAST_BOUNCE"}
        print(json.dumps(line))
    r.close()
    if dist:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--kernel", default="auto", choices=["auto", "mega", "wavefront", "server"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-jit", action="store_true", help="ahead-of-time kernel variants only (no per-scene NVRTC build)")
    ap.add_argument("--workload", default="cornell", choices=sorted(WORKLOADS), help="cornell = the headline; others are extra")
    args = ap.parse_args()
    global _workload
    _workload = args.workload
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
