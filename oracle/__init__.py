"""CPU oracle binding -- TEST INFRASTRUCTURE ONLY (see oracle/oracle.cpp).

Importable only from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs.  PARITY UNPINNED (no reference golden vectors exist).
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_f32p = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")


def build(force=False):
    """Compile liboracle.so / liboracle_ops.so with oracle/Makefile (g++ only)."""
    if force:
        subprocess.check_call(["make", "-C", _HERE, "clean"], stdout=subprocess.DEVNULL)
    subprocess.check_call(["make", "-C", _HERE, "-j2"], stdout=subprocess.DEVNULL)


def _opt(arr):
    return None if arr is None else arr.ctypes.data_as(ctypes.c_void_p)


class Oracle:
    """One loaded copy of the oracle library (``count_ops=True`` loads the op-counting build)."""

    def __init__(self, count_ops=False, ieee_div=False):
        """``ieee_div=True`` loads the build whose '/' is IEEE division and whose normalize() divides by the length
        (oracle/glsl.h SAIL_ORACLE_IEEE_DIV) -- for measuring the distance between the two arithmetic models only."""
        name = "liboracle_ieee.so" if ieee_div else ("liboracle_ops.so" if count_ops else "liboracle.so")
        path = os.path.join(_HERE, name)
        if not os.path.exists(path):
            build()
        self.lib = lib = ctypes.CDLL(path)
        lib.oracle_set_scene.argtypes = [_f32p, ctypes.c_int, _f32p, ctypes.c_int, _f32p, ctypes.c_int]
        lib.oracle_render_frame.restype = ctypes.c_uint64
        lib.oracle_render_frame.argtypes = [ctypes.c_int, ctypes.c_int, _f32p, _f32p, ctypes.c_float, ctypes.c_float,
                                            ctypes.c_int, _f32p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                            ctypes.c_void_p, ctypes.c_int]
        lib.oracle_trace_dump.argtypes = [ctypes.c_int, ctypes.c_int, _f32p, _f32p, ctypes.c_float, ctypes.c_int, _i32p,
                                          _f32p, _f32p, ctypes.c_void_p]
        lib.oracle_intersect_rays.argtypes = [ctypes.c_int, _f32p, _i32p, _f32p, ctypes.c_void_p, ctypes.c_void_p]
        lib.oracle_intersect_detail.argtypes = [_f32p, _f32p]
        lib.oracle_unit.argtypes = [ctypes.c_int, _f32p, _f32p]
        lib.oracle_random2.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_float, _f32p]
        lib.oracle_detmath.argtypes = [ctypes.c_int, ctypes.c_int, _f32p, ctypes.c_void_p, _f32p]
        lib.oracle_present.argtypes = [ctypes.c_int, ctypes.c_int, _f32p, _f32p, _f32p, ctypes.c_int, ctypes.c_float,
                                       _f32p, _f32p, ctypes.c_void_p, ctypes.c_void_p]

    def set_russian_roulette(self, start_bounce):
        """Opt-in extension (not in the reference): Russian roulette from bounce ``start_bounce`` on; 0 = off."""
        self.lib.oracle_set_russian_roulette(int(start_bounce))

    def set_scene(self, objects, texparams, lights):
        o = np.ascontiguousarray(objects, dtype=np.float32).reshape(-1)
        t = np.ascontiguousarray(texparams, dtype=np.float32).reshape(-1)
        l = np.ascontiguousarray(lights, dtype=np.float32).reshape(-1)
        pad = np.zeros(1, np.float32)
        rc = self.lib.oracle_set_scene(o if o.size else pad, o.size // 18, t if t.size else pad, t.size // 16,
                                       l if l.size else pad, l.size // 18)
        if rc:
            raise ValueError("oracle_set_scene failed")

    def render_frame(self, W, H, matrix, eye, textureWeight, time, accum, max_bounces=5, aovs=False, nthreads=0):
        """One progressive frame; ``accum`` (H,W,3 float32) is updated in place.  Returns
        dict(ops=..., normal, position, hit_id, depth) (AOVs only when aovs=True)."""
        matrix = np.ascontiguousarray(matrix, dtype=np.float32).reshape(16)
        eye = np.ascontiguousarray(eye, dtype=np.float32).reshape(3)
        assert accum.dtype == np.float32 and accum.shape == (H, W, 3) and accum.flags.c_contiguous
        out = {}
        if aovs:
            out["normal"] = np.zeros((H, W, 3), np.float32)
            out["position"] = np.zeros((H, W, 3), np.float32)
            out["hit_id"] = np.zeros((H, W), np.int32)
            out["depth"] = np.zeros((H, W), np.float32)
        out["ops"] = int(self.lib.oracle_render_frame(
            W, H, matrix, eye, float(textureWeight), float(time), int(max_bounces), accum.reshape(-1),
            _opt(out.get("normal")), _opt(out.get("position")), _opt(out.get("hit_id")), _opt(out.get("depth")),
            int(nthreads)))
        return out

    def trace_dump(self, W, H, matrix, eye, time, max_bounces=5):
        matrix = np.ascontiguousarray(matrix, dtype=np.float32).reshape(16)
        eye = np.ascontiguousarray(eye, dtype=np.float32).reshape(3)
        hit = np.zeros((H, W, max_bounces), np.int32)
        t = np.zeros((H, W, max_bounces), np.float32)
        rays = np.zeros((H, W, max_bounces, 6), np.float32)
        rad = np.zeros((H, W, 3), np.float32)
        self.lib.oracle_trace_dump(W, H, matrix, eye, float(time), int(max_bounces), hit.reshape(-1), t.reshape(-1),
                                   rays.reshape(-1), _opt(rad))
        return {"hit_id": hit, "t": t, "rays": rays, "radiance": rad}

    def intersect_rays(self, rays):
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 6)
        n = rays.shape[0]
        ids = np.zeros(n, np.int32)
        t = np.zeros(n, np.float32)
        normal = np.zeros((n, 3), np.float32)
        mat = np.zeros(n, np.int32)
        self.lib.oracle_intersect_rays(n, rays.reshape(-1), ids, t, _opt(normal), _opt(mat))
        return ids, t, normal, mat

    def intersect_detail(self, ray):
        out = np.zeros(24, np.float32)
        self.lib.oracle_intersect_detail(np.ascontiguousarray(ray, dtype=np.float32).reshape(6), out)
        return {"d": out[0], "hit": out[1:4], "normal": out[4:7], "dpdu": out[7:10], "sc": out[10:13],
                "emission": out[13:16], "into": bool(out[16]), "matCategory": int(out[17]), "matIndex": out[18],
                "index": int(out[19])}

    def random2(self, x, y, seed):
        out = np.zeros(2, np.float32)
        self.lib.oracle_random2(int(x), int(y), float(seed), out)
        return out

    def detmath(self, which, a, b=None):
        names = {"sin": 0, "cos": 1, "tan": 2, "atan": 3, "acos": 4, "asin": 5, "exp": 6, "log": 7, "atan2": 8, "pow": 9, "rcp": 12}
        a = np.ascontiguousarray(a, dtype=np.float32).reshape(-1)
        out = np.zeros_like(a)
        bb = None if b is None else np.ascontiguousarray(b, dtype=np.float32).reshape(-1)
        self.lib.oracle_detmath(names[which], a.size, a, _opt(bb), out)
        return out

    UNITS = {"frDielectric": (0, 1), "frConductor": (1, 3), "trowbridgeReitz_d": (2, 1), "trowbridgeReitz_sample_wh": (3, 3),
             "trowbridgeReitz_pdf": (4, 1), "getSurfaceColor": (5, 3), "cosineSampleHemisphere": (6, 3),
             "uniformSampleSphere": (7, 3), "concentricSampleDisk": (8, 2), "quadratic": (9, 3), "light_arm": (10, 3)}

    def unit(self, name, *args):
        """One shader function on scalar/vector arguments (flattened in order); see oracle_unit in oracle.cpp."""
        which, nout = self.UNITS[name]
        a = np.concatenate([np.asarray(x, np.float32).reshape(-1) for x in args]).astype(np.float32)
        a = np.concatenate([a, np.zeros(16, np.float32)])
        out = np.zeros(4, np.float32)
        if self.lib.oracle_unit(which, a, out) != 0:
            raise ValueError("oracle_unit(%s) rejected its arguments" % name)
        return out[:nout].copy() if nout > 1 else out[0]

    def present(self, color, normal, position, filt, as_u8=False):
        """Display pass.  ``filt`` = dict(filter_id, gamma_c, radius, weights) from sail_b200.filters.lower_filter."""
        H, W, _ = color.shape
        rgba = np.zeros((H, W, 4), np.float32)
        u8 = np.zeros((H, W, 4), np.uint8) if as_u8 else None
        self.lib.oracle_present(W, H, np.ascontiguousarray(color, np.float32).reshape(-1),
                                np.ascontiguousarray(normal, np.float32).reshape(-1),
                                np.ascontiguousarray(position, np.float32).reshape(-1), int(filt["filter_id"]),
                                float(filt["gamma_c"]), np.asarray(filt["radius"], np.float32),
                                np.asarray(filt["weights"], np.float32), _opt(rgba), _opt(u8))
        return (rgba, u8) if as_u8 else rgba
