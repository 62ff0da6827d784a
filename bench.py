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
        if s == 0:
            r.ctx.specialise_wait()
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


def _share_group_id(dist, rank):
    """The 128-byte rendezvous token of sail_group_join, made on rank 0 and carried to the others by torch.distributed --
    the only thing the bench uses torch.distributed for besides its barriers and the max over ranks."""
    from sail_b200 import capi
    box = [capi.group_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    return box[0]


def sustained_leg(local_rank, peak_tops, min_seconds=2.0):
    """BASELINE config 4 in full on one GPU -- Cornell box 3840x2160, 1024 spp as 32 launches of 32 spp, repeated until at
    least ``min_seconds`` of device time -- with its own clock record: the SM clock under seconds of sustained FP32 load."""
    global _workload
    import torch
    from sail_b200.renderer import Renderer
    keep, _workload = _workload, "config4"
    try:
        scene, W, H, bounces, _ = build_workload()
        ops = ops_per_path()
    finally:
        _workload = keep
    r = Renderer((W, H), device=local_rank)
    r.ctx.configure(max_bounces=bounces)
    r.update(scene)
    launches_per_pass = 1024 // SPP_PER_STEP
    batches = [r.tracer.frames(scene.mat, scene.eye, s * SPP_PER_STEP, SPP_PER_STEP) for s in range(launches_per_pass)]
    r.ctx.render_n(batches[0])
    r.ctx.specialise_wait()
    r.ctx.render_n(batches[1])
    r.ctx.synchronize()
    stream = torch.cuda.ExternalStream(r.ctx.stream(), device=local_rank)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler = ClockSampler(local_rank)
    sampler.start()
    passes, ms = 0, 0.0
    while ms < min_seconds * 1e3 and passes < 16:
        r.ctx.reset()
        with torch.cuda.stream(stream):
            e0.record(stream)
            for b in batches:
                r.ctx.render_n(b)
            e1.record(stream)
        r.ctx.synchronize()
        torch.cuda.synchronize()
        ms += e0.elapsed_time(e1)
        passes += 1
    clocks = sampler.stop()
    used = r.ctx.kernel_info()
    r.close()
    paths_s = passes * W * H * 1024 / (ms * 1e-3)
    return {"workload": WORKLOADS["config4"][1] + ", 1024 spp per pass, sustained", "value": paths_s / 1e6, "unit": UNIT,
            "seconds": ms * 1e-3, "passes": passes, "launches": passes * launches_per_pass, "kernel_used": used,
            "roofline_frac": paths_s * ops / 1e12 / peak_tops, "ops_per_path": ops, "clocks": clocks}


def interactive_leg(local_rank, frames=600):
    """BASELINE configs[0] the way the reference runs it (ui/ui.js:67-70): 512x512 Cornell box, ONE progressive sample per
    ``Renderer.render(scene)`` call, display pass and read-back of the RGBA8 canvas every call.  Wall clock around ``frames``
    calls, synchronous (each call returns its own image) and with the two-slot swapchain (``render_pipelined``); the third
    figure leaves the Python host work (jitter + inverse MVP per frame) outside: the same C-ABI calls on prepared frames."""
    from sail_b200 import configs
    from sail_b200.renderer import Renderer
    scene, W, H, _, bounces = configs.config1()
    r = Renderer((W, H), device=local_rank)
    r.ctx.configure(max_bounces=bounces)
    r.update(scene)
    r.render(scene)
    r.ctx.specialise_wait()
    for _ in range(20):
        r.render(scene)
    r.ctx.synchronize()
    t0 = time.perf_counter()
    for _ in range(frames):
        img = r.render(scene)
    r.ctx.synchronize()
    sync_s = time.perf_counter() - t0
    assert img.shape == (H, W, 4)
    for _ in range(4):
        r.render_pipelined(scene)
    r.flush()
    t0 = time.perf_counter()
    for _ in range(frames):
        r.render_pipelined(scene)
    img = r.flush()
    r.ctx.synchronize()
    pipe_s = time.perf_counter() - t0
    assert img.shape == (H, W, 4)
    prepared = [r.tracer.frames(scene.mat, scene.eye, scene.sampleCount + i, 1) for i in range(frames)]
    t0 = time.perf_counter()
    for fr in prepared:
        r.ctx.render_n(fr)
        img = r.ctx.present(as_u8=True, pinned=True)
    r.ctx.synchronize()
    cabi_s = time.perf_counter() - t0
    used = r.ctx.kernel_info()
    r.close()
    return {"workload": "BASELINE configs[0]: Cornell box 512x512, 1 spp per Renderer.render() call + display pass + RGBA8 read-back",
            "frames": frames, "frames_per_s": frames / sync_s, "ms_per_frame": sync_s / frames * 1e3,
            "frames_per_s_pipelined": frames / pipe_s, "frames_per_s_c_abi_prepared_frames": frames / cabi_s,
            "mpaths_per_s": frames * W * H / sync_s / 1e6, "d2h_bytes_per_frame": W * H * 4, "kernel_used": used,
            "note": "the reference paces this loop with requestAnimationFrame (display refresh, typically 60 Hz)"}


def run_ours(args, rank, world, local_rank):
    import torch

    from sail_b200 import capi
    from sail_b200.filters import lower_filter
    from sail_b200.generator import PluginParams
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
    tile = args.sharding == "tile" and world > 1
    shard = capi.SHARD_TILE if tile else capi.SHARD_SAMPLE
    # sample sharding: every rank traces SPP_PER_STEP frames per step (weak scaling); tile sharding: the N ranks share the
    # rows of SPP_PER_STEP frames (strong scaling)
    frames_per_step = SPP_PER_STEP if (tile or world == 1) else SPP_PER_STEP * world
    paths_per_step = W * H * frames_per_step  # whole job

    # One renderer per rank; with N > 1 it joins the group: the split of every sail_render_n(), the collective and the
    # combined image all live behind the C ABI (sail_group_join).
    group = (world, rank, _share_group_id(dist, rank)) if world > 1 else None
    r = Renderer((W, H), device=local_rank, group=group, shard=shard, band_rows=args.band_rows)
    ctx = r.ctx
    ctx.configure(max_bounces=bounces,
                  kernel={"wavefront": capi.KERNEL_WAVEFRONT, "mega": capi.KERNEL_MEGA, "auto": capi.KERNEL_AUTO, "queue": capi.KERNEL_QUEUE}[args.kernel],
                  specialise=capi.SPECIALISE_OFF if args.no_jit else capi.SPECIALISE_AUTO)
    r.update(scene)
    filters = []
    if args.sharding == "tile":  # BASELINE config 5: the filter sweep runs on the root after the gather (N = 1: same display work)
        for name in args.filters.split(","):
            pp = PluginParams(name)
            pp.addParam("r", "vec2(2.0,2.0)")
            pp.addParam("tau", "3.0")
            pp.addParam("alpha", "2.0")
            filters.append(lower_filter(pp))
    total_steps = warm + steps

    # ---------------- device-resident arm: the trace kernels alone, CUDA events on each rank's stream -----------------
    batches = [r.tracer.frames(scene.mat, scene.eye, s * frames_per_step, frames_per_step) for s in range(total_steps)]
    stream = torch.cuda.ExternalStream(ctx.stream(), device=local_rank)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    sampler = ClockSampler(local_rank)
    sampler.start()  # runs through the warm-up (same kernel, same load) and the timed region
    for s in range(warm):
        ctx.render_n(batches[s])
        if s == 0:
            ctx.specialise_wait()  # NVRTC works on another thread: the remaining warm-up steps already run the kernel that is timed
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

    # ---------------- the combine alone (device time on the root, warmed) -----------------
    combine_ms = None
    if dist:
        samples = []
        for i in range(6):
            ctx.render_n(batches[i % total_steps])
            img = ctx.read_accum()          # collective: ncclReduce to rank 0 (+ scale), then the read-back on rank 0
            if rank == 0 and i >= 2:
                samples.append(ctx.last_combine_ms())
        if rank == 0:
            combine_ms = float(np.median(samples))
            assert np.isfinite(img).all() and 0.0 < float(img.mean()) < 1.0, "combined image is not sane"

    # ---------------- end-to-end arm: Renderer API, host buffers in and out, combined image on rank 0 -----------------
    def present_sweep(pipelined):
        """One step's display work: with a filter sweep (config 5) every filter is applied to the same gathered image."""
        if not filters:
            return r.render_pipelined(scene, frames=frames_per_step) if pipelined else r.render(scene, frames=frames_per_step)
        r.trace(scene, frames_per_step)
        out = None
        for f in filters:
            ctx.configure(filter_id=f["filter_id"], filter_gamma_c=f["gamma_c"], filter_radius=f["radius"], filter_weights=f["weights"])
            out = ctx.present(as_u8=True, pinned=True)
        return out

    scene.sampleCount = 0
    e2e_steps = max(1, min(steps, 20))
    for _ in range(2):
        present_sweep(False)
    ctx.synchronize()
    if dist:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        # host: N x 32 x (jitter, inverse MVP) -> sail_frame[] -> H2D as kernel parameters; trace on every GPU; combine on the
        # root (N > 1); display filter; D2H of the RGBA8 image into page-locked host memory (synchronises on the root)
        img8 = present_sweep(False)
    ctx.synchronize()
    e2e_s = time.perf_counter() - t0
    assert rank != 0 or img8.shape == (H, W, 4)
    # the same steps with the display double-buffered (Renderer.render_pipelined: step i's image is collected by call i+1,
    # its device-to-host copy overlaps the tracing of step i+1; every copy still lies inside the timed region)
    e2e_pipe_s = None
    if not filters:
        for _ in range(2):
            present_sweep(True)
        r.flush()
        ctx.synchronize()
        if dist:
            dist.barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            img8 = present_sweep(True)
        img8 = r.flush()
        ctx.synchronize()
        e2e_pipe_s = time.perf_counter() - t0
        assert rank != 0 or img8.shape == (H, W, 4)
    group_info = ctx.group_info()[2]

    times = torch.tensor([dev_ms, e2e_s * 1e3, (e2e_pipe_s or e2e_s) * 1e3], dtype=torch.float64, device="cuda")
    tmin = times.clone()
    if dist:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
        dist.all_reduce(tmin, op=dist.ReduceOp.MIN)
    dev_ms, e2e_ms, e2e_pipe_ms = float(times[0]), float(times[1]), float(times[2])
    if rank == 0:
        pk = peaks()
        n_sm = torch.cuda.get_device_properties(local_rank).multi_processor_count
        value = steps * paths_per_step / (dev_ms * 1e-3) / 1e6
        per_gpu_paths_s = steps * paths_per_step / world / (dev_ms * 1e-3)
        ops = ops_per_path()
        peak = n_sm * 128 * pk["sm_max_mhz"] * 1e6 / 1e12
        achieved = per_gpu_paths_s * ops / 1e12
        launches_per_step = max(1, launches // max(1, steps))
        hbm_gbs = (W * H * 24 / (world if tile else 1)) * launches_per_step / (dev_ms / steps * 1e-3) / 1e9
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": warm,
            "ms_per_step": dev_ms / steps, "higher_is_better": True, "scaling": "strong" if tile else "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": WORKLOADS[_workload][1], "spp_per_step": SPP_PER_STEP, "paths_per_step": paths_per_step,
                       "kernel": args.kernel, "kernel_used": kernel_used, "arithmetic": "strict (bit-exact vs oracle)",
                       "sharding": ("interleaved %d-row bands mod N (tile)" % (args.band_rows or 8)) if tile else "sample index mod N",
                       "group": group_info,
                       "l2": "flushed between steps (256 MiB memset outside the per-step events)"},
            "spp_per_s": value * 1e6 / (W * H),
            "roofline": {"bound": "fp32_issue", "achieved": achieved, "peak": peak, "unit": "Tops/s (FP32 lane-ops)",
                         "frac": achieved / peak, "traffic": measured_traffic(), "ncu": measured_ncu(), "ops_per_path": ops,
                         "peak_def": "%d SMs x 128 lanes x %.0f MHz (%s sm_max_mhz)" % (n_sm, pk["sm_max_mhz"], pk["source"]),
                         "hbm": {"achieved": hbm_gbs, "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": hbm_gbs / pk["hbm_gbs"],
                                 "bytes_per_launch": W * H * 24}},
            "clocks": clocks,
            "e2e": {"value": e2e_steps * paths_per_step / (e2e_pipe_ms * 1e-3) / 1e6, "unit": UNIT,
                    "h2d_bytes_per_step": frames_per_step * 84, "d2h_bytes_per_step": W * H * 4 * max(1, len(filters)), "steps": e2e_steps,
                    "mode": ("Renderer.render_pipelined per step" if e2e_pipe_s is not None else "trace + one sail_present per filter per step") +
                            ": host frame construction, H2D of the frame uniforms, trace on every GPU, "
                            + ("combine of the %d GPUs' buffers on rank 0 INSIDE the timed region (%s), " % (world, group_info) if world > 1 else "")
                            + "display filter, D2H of the RGBA8 image into page-locked memory"
                            + ("; the display is double-buffered (step i's image is collected by call i+1, its copy overlaps the next trace), "
                               "every copy lies inside the timed region" if e2e_pipe_s is not None else ""),
                    "synchronous_value": e2e_steps * paths_per_step / (e2e_ms * 1e-3) / 1e6,
                    "synchronous_mode": "Renderer.render per step: the same work, the call returns its own image (waits for the copy)"},
            "gpu_launches": int(launches),
        }
        if world > 1:
            line["trace_ms_per_step_min_over_ranks"] = float(tmin[0]) / steps
            line["trace_ms_per_step_max_over_ranks"] = dev_ms / steps
            line["combine"] = {"ms": combine_ms, "bytes": W * H * 12, "gbs": W * H * 12 / (combine_ms * 1e-3) / 1e9 if combine_ms else None,
                               "how": group_info, "note": "device time on rank 0 from the end of its own trace to the combined buffer, median of 4 warmed calls"}
            if filters:
                line["config"]["filters"] = args.filters
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_baseline()
        if world == 1 and not args.no_cpu and _workload == "cornell":
            line["other_workloads"] = [other_workload(w, local_rank, peak) for w in ("config2", "config3", "config5")]
            line["interactive"] = interactive_leg(local_rank)
            line["other_workloads"].append(sustained_leg(local_rank, peak))
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
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--kernel", default="auto", choices=["auto", "mega", "wavefront", "queue"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-jit", action="store_true", help="ahead-of-time kernel variants only (no per-scene NVRTC build)")
    ap.add_argument("--workload", default="cornell", choices=sorted(WORKLOADS), help="cornell = the headline; others are extra")
    ap.add_argument("--sharding", default="sample", choices=["sample", "tile"], help="N > 1: sample index mod N (config 4) or row bands (config 5)")
    ap.add_argument("--band-rows", type=int, default=0, help="tile sharding: rows per interleaved band (multiple of 8; 0 = 8)")
    ap.add_argument("--filters", default="box,sinc,gaussian", help="tile sharding: display filters applied to the gathered image each step")
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
