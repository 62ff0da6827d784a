"""Renderer / Tracer: the host-side mirror of the reference's render orchestration
(/root/reference/src/core/renderer.js:8-73, src/core/tracer.js:8-101), driving libsail_b200.so.

Same names, argument meaning and call order as the reference: ``renderer.update(scene)`` commits the
scene (plugin config + the three wire-format tables), ``renderer.render(scene)`` traces one sample per
pixel and runs the display filter, ``renderer.updateObjects(scene)`` re-uploads the objects table.
Differences forced by the platform: the canvas is a ``(width, height)`` pair, the displayed image is
returned as an array instead of being drawn, and the wall-clock seed / ``Math.random()`` jitter of
tracer.js:94-98 come from an explicit :class:`FrameSchedule` so frames are reproducible.
"""
import numpy as np

from . import capi
from .filters import lower_filter
from .scene import plugin_masks
from .schedule import FrameSchedule, frame_matrices, frame_matrix
from .wire import serialize_objects, serialize_scene


class Tracer:
    """tracer.js:8-101."""

    def __init__(self, ctx, schedule=None):
        self.ctx = ctx
        self.schedule = schedule or FrameSchedule()
        self.n = self.tn = self.ln = 0

    def updateObjects(self, scene):
        """tracer.js:25-40."""
        objects = serialize_objects(scene)
        self.n = objects.shape[0]
        self.ctx.upload_objects(objects)

    def update(self, scene):
        """tracer.js:42-90: setProgram(new TraceShader(scene.tracerConfig())) + table upload."""
        masks = plugin_masks(scene.tracerConfig())
        self.ctx.configure(**masks)
        objects, texparams, lights = serialize_scene(scene)
        self.n, self.tn, self.ln = objects.shape[0], texparams.shape[0], lights.shape[0]
        self.ctx.upload_scene(objects, texparams, lights)

    def frame(self, modelviewProjection, eye, sampleCount):
        """The uniforms of tracer.js:92-98 for progressive sample ``sampleCount``."""
        k = int(sampleCount)
        m = frame_matrix(modelviewProjection, self.schedule.jitter_unit(k), self.ctx.width, self.ctx.height)
        return capi.make_frame(m, eye, k / (k + 1.0), self.schedule.time(k))

    def frames(self, modelviewProjection, eye, sampleCount, count):
        """``count`` consecutive frames starting at ``sampleCount`` as one sail_frame[] (same values as ``frame``)."""
        k0 = int(sampleCount)
        m = frame_matrices(modelviewProjection, self.schedule.jitter_units(k0, count), self.ctx.width, self.ctx.height)
        k = np.arange(k0, k0 + count, dtype=np.float64)
        return capi.frames_from_arrays(m, eye, k / (k + 1.0), self.schedule.times(k0, count))

    def render(self, modelviewProjection, eye, sampleCount):
        """tracer.js:92-101."""
        f = self.frame(modelviewProjection, eye, sampleCount)
        self.ctx.render_n([f])


class Renderer:
    """renderer.js:8-73.  ``canvas`` is ``(width, height)`` or an object with ``width``/``height``.

    Multi-GPU (the reference has one GL context): ``devices=[0, 1, ...]`` makes one renderer over several
    GPUs of this process, ``group=(nranks, rank, group_id)`` makes this renderer one member of a group with one GPU per
    process (every member then issues the same calls; images arrive on rank 0).  ``shard`` = capi.SHARD_SAMPLE or
    capi.SHARD_TILE; the split, the collective and the combined image all live behind the C ABI."""

    def __init__(self, canvas, device=0, rows=None, schedule=None, devices=None, shard=capi.SHARD_SAMPLE, band_rows=0,
                 combine=capi.COMBINE_AUTO, group=None):
        if isinstance(canvas, (tuple, list)):
            width, height = canvas
        elif isinstance(canvas, dict):
            width, height = canvas["width"], canvas["height"]
        else:
            width, height = canvas.width, canvas.height
        self.ctx = capi.Context(width, height, device=device, rows=rows, devices=devices, shard=shard, band_rows=band_rows,
                                combine=combine)
        if group is not None:
            nranks, rank, group_id = group
            self.ctx.group_join(nranks, rank, group_id, shard=shard, band_rows=band_rows, combine=combine)
        self.tracer = Tracer(self.ctx, schedule)
        self._filter = None

    def close(self):
        self.ctx.close()

    def updateObjects(self, scene):
        self.tracer.updateObjects(scene)

    def update(self, scene):
        """renderer.js:45-52."""
        f = lower_filter(scene.rendererConfig()["filter"])
        self._filter = f
        self.ctx.configure(filter_id=f["filter_id"], filter_gamma_c=f["gamma_c"], filter_radius=f["radius"],
                           filter_weights=f["weights"])
        self.tracer.update(scene)

    def trace(self, scene, frames=1):
        """``frames`` progressive samples without the display pass (batched through sail_render_n)."""
        if scene.moving:
            scene.sampleCount = 0
            self.updateObjects(scene)
        if scene.sampleCount == 0:
            self.ctx.reset()
        fr = self.tracer.frames(scene.mat, scene.eye, scene.sampleCount, frames)
        scene.sampleCount += frames
        self.ctx.render_n(fr)

    def render(self, scene, as_u8=True, frames=1):
        """renderer.js:54-72: one progressive sample (or ``frames`` of them), then the display pass; returns the
        displayed image (H x W x 4, row 0 = bottom) in the renderer's page-locked "canvas" buffer, which the next
        ``render`` overwrites.  The selection wireframe (renderer.js:65-71) is editor UI, out of scope."""
        self.trace(scene, frames)
        return self.ctx.present(as_u8=as_u8, pinned=True)

    def render_pipelined(self, scene, frames=1):
        """``render`` with a swapchain of two: the display pass and the device-to-host copy of THIS call's image are only
        queued (sail_present_begin), and the image of the PREVIOUS call -- complete by now, its copy overlapped this call's
        tracing -- is returned (``None`` on the first call).  ``flush()`` returns the last image.  This is the pacing of the
        reference's loop: ``renderer.render(scene)`` per requestAnimationFrame, the canvas reaching the screen a frame later
        (ui/ui.js:67-70)."""
        self.trace(scene, frames)
        prev = self.flush()
        self._slot = 1 - getattr(self, "_slot", 1)
        self._inflight = self.ctx.present_begin(self._slot)  # None on the non-root members of a group
        self._pending = True
        return prev

    def flush(self):
        """Wait for the image queued by the last ``render_pipelined`` and return it (``None`` if there is none)."""
        if not getattr(self, "_pending", False):
            return None
        self.ctx.present_end()
        self._pending = False
        buf, self._inflight = self._inflight, None
        return buf

    def save(self, path, img=None):
        """Write the displayed image (``img`` or a fresh display pass of the current accumulation) as .png or .ppm --
        the headless stand-in for the canvas of renderer.js:63."""
        from . import image
        if img is None:
            img = self.ctx.present(as_u8=True, pinned=True)
        (image.write_ppm if str(path).lower().endswith(".ppm") else image.write_png)(path, img)

    def readPixels(self, which=capi.AOV_COLOR):
        return self.ctx.read_aov(which)
