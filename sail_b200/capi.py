"""ctypes binding of ``libsail_b200.so`` (include/sail_b200.h).

This is the Python twin of the N-API shim a Node.js host would use: it only marshals
pointers and sizes across the C ABI.  There is no fallback -- if the library is missing or no sm_100
device is present the calls raise.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SAIL_B200_LIB") or os.path.join(_HERE, "libsail_b200.so")  # SAIL_B200_LIB: A/B runs of two builds

MAX_FRAMES_PER_CALL = 32
AOV_COLOR, AOV_NORMAL, AOV_POSITION, AOV_HIT_ID, AOV_DEPTH = range(5)
ACCUM_MEAN, ACCUM_SUM = 0, 1
KERNEL_MEGA, KERNEL_WAVEFRONT, KERNEL_AUTO, KERNEL_QUEUE = 0, 1, 2, 3
SPECIALISE_AUTO, SPECIALISE_OFF, SPECIALISE_REQUIRE = 0, 1, 2
ERR_CUDA, ERR_INVALID, ERR_STATE, ERR_NOMEM = -1, -2, -3, -4
SHARD_NONE, SHARD_SAMPLE, SHARD_TILE = 0, 1, 2
COMBINE_AUTO, COMBINE_NCCL, COMBINE_P2P = 0, 1, 2
MAX_DEVICES = 8
GROUP_ID_BYTES = 128
ABI_VERSION = 2
NOALLOC_MAX_RAYS = 65536


class SailError(RuntimeError):
    def __init__(self, code, message):
        super().__init__("sail_b200 error %d: %s" % (code, message))
        self.code = code


class CreateInfo(ctypes.Structure):
    _fields_ = [("width", ctypes.c_int32), ("height", ctypes.c_int32), ("device", ctypes.c_int32),
                ("row_begin", ctypes.c_int32), ("row_end", ctypes.c_int32), ("ndev", ctypes.c_int32),
                ("shard_mode", ctypes.c_int32), ("band_rows", ctypes.c_int32), ("device_ids", ctypes.c_int32 * MAX_DEVICES),
                ("combine", ctypes.c_int32), ("reserved", ctypes.c_int32 * 3)]


class Config(ctypes.Structure):
    _fields_ = [("shape_mask", ctypes.c_uint32), ("material_mask", ctypes.c_uint32), ("texture_mask", ctypes.c_uint32),
                ("light_mask", ctypes.c_uint32), ("trace_id", ctypes.c_int32), ("max_bounces", ctypes.c_int32),
                ("accum_mode", ctypes.c_int32), ("kernel", ctypes.c_int32), ("filter_id", ctypes.c_int32),
                ("filter_gamma_c", ctypes.c_float), ("filter_radius", ctypes.c_float * 2),
                ("filter_weights", ctypes.c_float * 16), ("specialise", ctypes.c_int32), ("rr_start_bounce", ctypes.c_int32),
                ("reserved", ctypes.c_int32 * 2)]


class Frame(ctypes.Structure):
    _fields_ = [("matrix", ctypes.c_float * 16), ("eye", ctypes.c_float * 3), ("textureWeight", ctypes.c_float),
                ("timeSinceStart", ctypes.c_float)]


#: every symbol include/sail_b200.h declares: name -> (restype, argtypes)
_P = ctypes.c_void_p
_SIGNATURES = {
    "sail_version": (ctypes.c_char_p, []),
    "sail_abi_version": (ctypes.c_int, []),
    "sail_device_count": (ctypes.c_int, []),
    "sail_create": (ctypes.c_int, [ctypes.POINTER(CreateInfo), ctypes.POINTER(_P)]),
    "sail_destroy": (None, [_P]),
    "sail_group_id": (ctypes.c_int, [_P]),
    "sail_group_join": (ctypes.c_int, [_P, ctypes.c_int32, ctypes.c_int32, _P, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32]),
    "sail_group_info": (ctypes.c_int, [_P, ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_int32), ctypes.c_char_p, ctypes.c_int32]),
    "sail_last_combine_ms": (ctypes.c_int, [_P, ctypes.POINTER(ctypes.c_float)]),
    "sail_specialise_wait": (ctypes.c_int, [_P]),
    "sail_last_error": (ctypes.c_char_p, [_P]),
    "sail_default_config": (None, [ctypes.POINTER(Config)]),
    "sail_configure": (ctypes.c_int, [_P, ctypes.POINTER(Config)]),
    "sail_upload_scene": (ctypes.c_int, [_P, _P, ctypes.c_int32, _P, ctypes.c_int32, _P, ctypes.c_int32]),
    "sail_upload_objects": (ctypes.c_int, [_P, _P, ctypes.c_int32]),
    "sail_scene_features": (ctypes.c_int, [_P, ctypes.c_int32, _P, ctypes.c_int32, _P, ctypes.c_int32, ctypes.c_char_p, ctypes.c_int32]),
    "sail_set_row_bands": (ctypes.c_int, [_P, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32]),
    "sail_reset": (ctypes.c_int, [_P]),
    "sail_render": (ctypes.c_int, [_P, _P, _P, ctypes.c_float, ctypes.c_float]),
    "sail_render_n": (ctypes.c_int, [_P, ctypes.POINTER(Frame), ctypes.c_int32]),
    "sail_synchronize": (ctypes.c_int, [_P]),
    "sail_read_aov": (ctypes.c_int, [_P, ctypes.c_int32, _P, ctypes.c_size_t]),
    "sail_read_accum": (ctypes.c_int, [_P, _P, ctypes.c_size_t]),
    "sail_write_accum": (ctypes.c_int, [_P, _P, ctypes.c_size_t]),
    "sail_scale_accum": (ctypes.c_int, [_P, ctypes.c_float]),
    "sail_present": (ctypes.c_int, [_P, _P, _P]),
    "sail_present_begin": (ctypes.c_int, [_P, _P]),
    "sail_present_end": (ctypes.c_int, [_P]),
    "sail_device_ptr": (_P, [_P, ctypes.c_int32]),
    "sail_stream": (_P, [_P]),
    "sail_intersect_rays": (ctypes.c_int, [_P, _P, ctypes.c_int32, _P, _P, _P, _P]),
    "sail_intersect_rays_noalloc": (ctypes.c_int, [_P, _P, ctypes.c_int32, _P, _P, _P, _P]),
    "sail_trace_dump": (ctypes.c_int, [_P, ctypes.POINTER(Frame), _P, _P, _P, _P]),
    "sail_eval_math": (ctypes.c_int, [_P, ctypes.c_int32, _P, _P, _P, ctypes.c_int32]),
    "sail_selftest_math": (ctypes.c_int, [_P, ctypes.POINTER(ctypes.c_uint64)]),
    "sail_launch_count": (ctypes.c_uint64, [_P]),
    "sail_last_render_ms": (ctypes.c_int, [_P, ctypes.POINTER(ctypes.c_float)]),
    "sail_kernel_info": (ctypes.c_int, [_P, ctypes.c_char_p, ctypes.c_int32]),
    "sail_jit_compile_check": (ctypes.c_int, [_P, ctypes.c_int32, _P, ctypes.c_int32, _P, ctypes.c_int32, ctypes.c_int32,
                                              ctypes.c_char_p, ctypes.c_int32]),
    "sail_alloc_host": (ctypes.c_void_p, [ctypes.c_uint64]),
    "sail_free_host": (None, [ctypes.c_void_p]),
}
EXPORTED_SYMBOLS = tuple(_SIGNATURES)

_lib = None


def build(verbose=False):
    """Compile libsail_b200.so in-tree with nvcc for sm_100a (sail_b200/csrc/Makefile)."""
    out = None if verbose else subprocess.DEVNULL
    subprocess.check_call(["make", "-C", os.path.join(_HERE, "csrc")], stdout=out)


def load():
    """Load the shared library (never builds implicitly, never falls back)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("%s is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(there is no CPU fallback)" % LIB_PATH)
        lib = ctypes.CDLL(LIB_PATH)
        for name, (restype, argtypes) in _SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = restype
            fn.argtypes = argtypes
        _lib = lib
    return _lib


def _f32(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.float32)
    return a if shape is None else a.reshape(shape)


def _ptr(a):
    return None if a is None else a.ctypes.data_as(_P)


def make_frame(matrix, eye, textureWeight, timeSinceStart):
    f = Frame()
    f.matrix[:] = [float(v) for v in np.asarray(matrix, dtype=np.float32).reshape(16)]
    f.eye[:] = [float(v) for v in np.asarray(eye, dtype=np.float32).reshape(3)]
    f.textureWeight = float(textureWeight)
    f.timeSinceStart = float(timeSinceStart)
    return f


class Context:
    """One ``sail_ctx``: one GPU (a device, a stream and the full-image device buffers) or, with ``devices=[...]``,
    one context sharded over several GPUs of this process (``shard`` = SHARD_SAMPLE / SHARD_TILE)."""

    def __init__(self, width, height, device=0, rows=None, devices=None, shard=SHARD_SAMPLE, band_rows=0, combine=COMBINE_AUTO):
        self.lib = load()
        self.width, self.height = int(width), int(height)
        info = CreateInfo(self.width, self.height, int(device), 0, 0)
        if rows is not None:
            info.row_begin, info.row_end = int(rows[0]), int(rows[1])
        if devices is not None and len(devices) > 1:
            if len(devices) > MAX_DEVICES:
                raise SailError(ERR_INVALID, "at most %d devices" % MAX_DEVICES)
            info.ndev = len(devices)
            info.device_ids[:len(devices)] = [int(d) for d in devices]
            info.shard_mode, info.band_rows, info.combine = int(shard), int(band_rows), int(combine)
        self.is_root = True
        self._sample_members = len(devices) if (devices is not None and len(devices) > 1 and shard == SHARD_SAMPLE) else 0
        handle = _P()
        self._pinned = {}
        rc = self.lib.sail_create(ctypes.byref(info), ctypes.byref(handle))
        self._h = handle
        if rc != 0:
            msg = self.lib.sail_last_error(handle).decode() if handle else "sail_create failed"
            if handle:
                self.lib.sail_destroy(handle)
            self._h = None
            raise SailError(rc, msg)
        self.config = Config()
        self.lib.sail_default_config(ctypes.byref(self.config))

    def close(self):
        if getattr(self, "_h", None):
            for p, _ in getattr(self, "_pinned", {}).values():
                self.lib.sail_free_host(p)
            self._pinned = {}
            self.lib.sail_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _check(self, rc):
        if rc != 0:
            raise SailError(rc, self.lib.sail_last_error(self._h).decode())

    # ---- configuration / upload -------------------------------------------------------------------------
    def configure(self, **kw):
        """Update fields of the sail_config (max_bounces, accum_mode, kernel, filter_id, filter_gamma_c,
        filter_radius, filter_weights, *_mask) and apply it."""
        for k, v in kw.items():
            if k in ("filter_radius", "filter_weights"):
                arr = getattr(self.config, k)
                arr[:] = [float(x) for x in v]
            else:
                if not hasattr(self.config, k):
                    raise AttributeError("sail_config has no field %r" % k)
                setattr(self.config, k, v)
        self._check(self.lib.sail_configure(self._h, ctypes.byref(self.config)))

    def upload_scene(self, objects, texparams, lights):
        o, t, l = _f32(objects).reshape(-1), _f32(texparams).reshape(-1), _f32(lights).reshape(-1)
        if o.size % 18 or t.size % 16 or l.size % 18:
            raise SailError(ERR_INVALID, "table sizes must be multiples of 18 / 16 / 18 floats")
        self._check(self.lib.sail_upload_scene(self._h, _ptr(o) if o.size else None, o.size // 18,
                                               _ptr(t) if t.size else None, t.size // 16,
                                               _ptr(l) if l.size else None, l.size // 18))

    def upload_objects(self, objects):
        o = _f32(objects).reshape(-1)
        self._check(self.lib.sail_upload_objects(self._h, _ptr(o) if o.size else None, o.size // 18))

    def reset(self):
        self._check(self.lib.sail_reset(self._h))

    def set_row_bands(self, band_rows, band_stride, band_first):
        self._check(self.lib.sail_set_row_bands(self._h, int(band_rows), int(band_stride), int(band_first)))

    # ---- rendering ------------------------------------------------------------------------------------
    def render(self, matrix, eye, textureWeight, timeSinceStart):
        m, e = _f32(matrix, 16), _f32(eye, 3)
        self._check(self.lib.sail_render(self._h, _ptr(m), _ptr(e), float(textureWeight), float(timeSinceStart)))

    def render_n(self, frames):
        """``frames``: a ctypes array of Frame (see ``frame_array``) or a list of Frame."""
        if isinstance(frames, (list, tuple)):
            frames = (Frame * len(frames))(*frames)
        count = len(frames)
        per_call = MAX_FRAMES_PER_CALL * (self._sample_members or 1)
        for start in range(0, count, per_call):
            n = min(per_call, count - start)
            chunk = ctypes.cast(ctypes.byref(frames, start * ctypes.sizeof(Frame)), ctypes.POINTER(Frame))
            self._check(self.lib.sail_render_n(self._h, chunk, n))

    def synchronize(self):
        self._check(self.lib.sail_synchronize(self._h))

    def specialise_wait(self):
        """Block until the scene-specialised kernel is in place; False when the context stays on the ahead-of-time variant."""
        rc = self.lib.sail_specialise_wait(self._h)
        if rc < 0:
            self._check(rc)
        return rc == 0

    # ---- groups: one GPU per process (torchrun-style) ---------------------------------------------------------
    def group_join(self, nranks, rank, group_id, shard=SHARD_SAMPLE, band_rows=0, combine=COMBINE_AUTO):
        """Collective: become member ``rank`` of ``nranks`` (``group_id`` = the 128 bytes of ``capi.group_id()`` made on
        one rank and carried to the others by the host)."""
        buf = (ctypes.c_uint8 * GROUP_ID_BYTES).from_buffer_copy(bytes(group_id))
        self._check(self.lib.sail_group_join(self._h, int(nranks), int(rank), buf, int(shard), int(band_rows), int(combine)))
        self.is_root = rank == 0
        self._sample_members = int(nranks) if shard == SHARD_SAMPLE else 0

    def group_info(self):
        n, first = ctypes.c_int32(), ctypes.c_int32()
        buf = ctypes.create_string_buffer(200)
        self._check(self.lib.sail_group_info(self._h, ctypes.byref(n), ctypes.byref(first), buf, 200))
        return int(n.value), int(first.value), buf.value.decode()

    def last_combine_ms(self):
        ms = ctypes.c_float()
        self._check(self.lib.sail_last_combine_ms(self._h, ctypes.byref(ms)))
        return ms.value

    def last_render_ms(self):
        ms = ctypes.c_float()
        self._check(self.lib.sail_last_render_ms(self._h, ctypes.byref(ms)))
        return ms.value

    def kernel_info(self):
        """Which trace kernel the last render used (scene-specialised at run time, or an ahead-of-time variant and why)."""
        buf = ctypes.create_string_buffer(400)
        self._check(min(0, self.lib.sail_kernel_info(self._h, buf, 400)))
        return buf.value.decode()

    def launch_count(self):
        return int(self.lib.sail_launch_count(self._h))

    # ---- outputs -------------------------------------------------------------------------------------
    def read_aov(self, which, out=None):
        H, W = self.height, self.width
        if which in (AOV_COLOR, AOV_NORMAL, AOV_POSITION):
            shape, dtype = (H, W, 3), np.float32
        elif which == AOV_HIT_ID:
            shape, dtype = (H, W), np.int32
        elif which == AOV_DEPTH:
            shape, dtype = (H, W), np.float32
        else:
            raise SailError(ERR_INVALID, "unknown AOV %r" % (which,))
        if not self.is_root:  # member of a multi-process group: takes part in the combine, the image arrives on rank 0
            self._check(self.lib.sail_read_aov(self._h, which, None, 0))
            return None
        if out is None:
            out = np.empty(shape, dtype)
        assert out.shape == shape and out.dtype == dtype and out.flags.c_contiguous
        self._check(self.lib.sail_read_aov(self._h, which, _ptr(out), out.nbytes))
        return out

    def read_accum(self, out=None):
        return self.read_aov(AOV_COLOR, out)

    def write_accum(self, rgb):
        a = _f32(rgb, (self.height, self.width, 3))
        self._check(self.lib.sail_write_accum(self._h, _ptr(a), a.nbytes))

    def scale_accum(self, s):
        self._check(self.lib.sail_scale_accum(self._h, float(s)))

    def host_buffer(self, shape, dtype):
        """A numpy array over page-locked memory from sail_alloc_host, cached per (shape, dtype) and freed with
        the context: the reusable destination of ``present(..., pinned=True)`` (the canvas of renderer.js:63)."""
        key = (tuple(shape), np.dtype(dtype).str)
        if key not in self._pinned:
            nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
            p = self.lib.sail_alloc_host(nbytes)
            if not p:
                raise SailError(ERR_NOMEM, "sail_alloc_host(%d) failed" % nbytes)
            arr = np.frombuffer((ctypes.c_uint8 * nbytes).from_address(p), dtype=dtype).reshape(shape)
            self._pinned[key] = (p, arr)
        return self._pinned[key][1]

    def present(self, as_u8=False, both=False, pinned=False):
        """Display pass + read-back.  ``pinned=True`` returns views of context-owned page-locked buffers that the
        next ``present`` overwrites (like the canvas); the default returns fresh arrays."""
        H, W = self.height, self.width
        if not self.is_root:
            self._check(self.lib.sail_present(self._h, None, None))
            return (None, None) if both else None
        alloc = self.host_buffer if pinned else (lambda shape, dtype: np.empty(shape, dtype))
        f32 = alloc((H, W, 4), np.float32) if (both or not as_u8) else None
        u8 = alloc((H, W, 4), np.uint8) if (both or as_u8) else None
        self._check(self.lib.sail_present(self._h, _ptr(f32), _ptr(u8)))
        if both:
            return f32, u8
        return u8 if as_u8 else f32

    def present_begin(self, slot):
        """Queue the display pass and start the asynchronous copy of the RGBA8 image into page-locked buffer ``slot``
        (0 or 1); returns that buffer, valid after ``present_end()``."""
        H, W = self.height, self.width
        if not self.is_root:
            self._check(self.lib.sail_present_begin(self._h, None))
            return None
        buf = self.host_buffer((2, H, W, 4), np.uint8)[slot]
        self._check(self.lib.sail_present_begin(self._h, buf.ctypes.data_as(ctypes.c_void_p)))
        return buf

    def present_end(self):
        self._check(self.lib.sail_present_end(self._h))

    def device_ptr(self, which):
        return self.lib.sail_device_ptr(self._h, which)

    def stream(self):
        return self.lib.sail_stream(self._h)

    # ---- diagnostics ------------------------------------------------------------------------------------
    def intersect_rays(self, rays):
        r = _f32(rays).reshape(-1, 6)
        n = r.shape[0]
        ids, t = np.empty(n, np.int32), np.empty(n, np.float32)
        normal, mat = np.empty((n, 3), np.float32), np.empty(n, np.int32)
        self._check(self.lib.sail_intersect_rays(self._h, _ptr(r), n, _ptr(ids), _ptr(t), _ptr(normal), _ptr(mat)))
        return ids, t, normal, mat

    def intersect_rays_noalloc(self, rays):
        """``intersect_rays`` through sail_intersect_rays_noalloc: fixed in-library scratch, at most NOALLOC_MAX_RAYS rays."""
        r = _f32(rays).reshape(-1, 6)
        n = r.shape[0]
        ids, t = np.empty(n, np.int32), np.empty(n, np.float32)
        normal, mat = np.empty((n, 3), np.float32), np.empty(n, np.int32)
        self._check(self.lib.sail_intersect_rays_noalloc(self._h, _ptr(r), n, _ptr(ids), _ptr(t), _ptr(normal), _ptr(mat)))
        return ids, t, normal, mat

    def trace_dump(self, frame):
        H, W, B = self.height, self.width, int(self.config.max_bounces)
        hit, t = np.empty((H, W, B), np.int32), np.empty((H, W, B), np.float32)
        rays, rad = np.empty((H, W, B, 6), np.float32), np.empty((H, W, 3), np.float32)
        self._check(self.lib.sail_trace_dump(self._h, ctypes.byref(frame), _ptr(hit), _ptr(t), _ptr(rays), _ptr(rad)))
        return {"hit_id": hit, "t": t, "rays": rays, "radiance": rad}

    def eval_math(self, which, a, b=None):
        names = {"sin": 0, "cos": 1, "tan": 2, "atan": 3, "acos": 4, "asin": 5, "exp": 6, "log": 7, "atan2": 8,
                 "pow": 9, "div": 10, "sqrt": 11, "rcp": 12}
        a = _f32(a).reshape(-1)
        bb = None if b is None else _f32(b).reshape(-1)
        out = np.empty_like(a)
        self._check(self.lib.sail_eval_math(self._h, names[which], _ptr(a), _ptr(bb), _ptr(out), a.size))
        return out


def _selftest_math(self):
    n = ctypes.c_uint64()
    self._check(self.lib.sail_selftest_math(self._h, ctypes.byref(n)))
    return int(n.value)


Context.selftest_math = _selftest_math


def scene_features(objects, texparams, lights):
    """Validate tables without a device; returns the feature bits or raises SailError."""
    o, t, l = _f32(objects).reshape(-1), _f32(texparams).reshape(-1), _f32(lights).reshape(-1)
    err = ctypes.create_string_buffer(256)
    rc = load().sail_scene_features(_ptr(o) if o.size else None, o.size // 18, _ptr(t) if t.size else None, t.size // 16,
                                    _ptr(l) if l.size else None, l.size // 18, err, 256)
    if rc < 0:
        raise SailError(rc, err.value.decode())
    return rc


def frame_array(frames):
    return (Frame * len(frames))(*frames)


def frames_from_arrays(matrices, eye, textureWeights, times):
    """A ``sail_frame[n]`` from (n, 16) column-major matrices, one eye, (n,) weights and (n,) times, filled in one
    numpy pass (a sail_frame is 21 consecutive floats: include/sail_b200.h)."""
    m = np.asarray(matrices, dtype=np.float32).reshape(-1, 16)
    n = m.shape[0]
    rows = np.empty((n, 21), dtype=np.float32)
    rows[:, :16] = m
    rows[:, 16:19] = np.asarray(eye, dtype=np.float32).reshape(3)
    rows[:, 19] = np.asarray(textureWeights, dtype=np.float32)
    rows[:, 20] = np.asarray(times, dtype=np.float32)
    out = (Frame * n)()
    ctypes.memmove(out, rows.ctypes.data, rows.nbytes)
    return out


def jit_compile_check(objects, texparams, lights, kernel=KERNEL_MEGA):
    """Compile (not load) the scene-specialised kernel of these tables; works without a GPU.  Returns the description."""
    o, t, l = _f32(objects).reshape(-1), _f32(texparams).reshape(-1), _f32(lights).reshape(-1)
    buf = ctypes.create_string_buffer(600)
    rc = load().sail_jit_compile_check(_ptr(o) if o.size else None, o.size // 18, _ptr(t) if t.size else None, t.size // 16,
                                       _ptr(l) if l.size else None, l.size // 18, int(kernel), buf, 600)
    if rc != 0:
        raise SailError(rc, buf.value.decode(errors="replace"))
    return buf.value.decode()


def device_count():
    return int(load().sail_device_count())


def group_id():
    """The 128-byte rendezvous token of a multi-process group (sail_group_id; needs NCCL)."""
    buf = (ctypes.c_uint8 * GROUP_ID_BYTES)()
    rc = load().sail_group_id(buf)
    if rc != 0:
        raise SailError(rc, "sail_group_id failed (NCCL not available?)")
    return bytes(buf)
