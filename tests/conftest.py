import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


# SAIL_SPECIALISE_AUTO compiles the scene's kernel on a worker thread and serves the first frames with the ahead-of-time
# variant; the suite is about the specialised kernels as much as about those, so here AUTO waits for the compiler
# (tests of the asynchronous hand-over remove the variable themselves).
os.environ.setdefault("SAIL_B200_JIT_SYNC", "1")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    """The CPU oracle (test infrastructure only; oracle/oracle.cpp)."""
    import oracle as oracle_pkg
    oracle_pkg.build()
    return oracle_pkg.Oracle()


def scene_tables(cfg, width, height, **kw):
    """(scene, objects, texparams, lights, max_bounces) of BASELINE config ``cfg`` at a test size."""
    from sail_b200 import configs, wire
    scene, W, H, frames, bounces = configs.CONFIGS[cfg](width, height, **kw)
    o, t, l = wire.serialize_scene(scene)
    return scene, o, t, l, bounces


def frame_uniforms(scene, k, W, H, schedule=None):
    from sail_b200 import schedule as sch
    s = schedule or sch.FrameSchedule()
    m = sch.frame_matrix(scene.mat, s.jitter_unit(k), W, H)
    return m, np.asarray(scene.eye, np.float32), k / (k + 1.0), s.time(k)


def _canon(a):
    """float32 bits with every NaN mapped to one pattern (x86 and sm_100a generate different default NaNs)."""
    a = np.ascontiguousarray(a, np.float32)
    bits = a.view(np.uint32).copy()
    bits[np.isnan(a)] = 0x7FC00000
    return bits


def bits_equal(a, b):
    """Bit-exact float comparison (signed zeros and infinities included; NaN == NaN)."""
    a = np.ascontiguousarray(a, np.float32)
    b = np.ascontiguousarray(b, np.float32)
    return a.shape == b.shape and np.array_equal(_canon(a), _canon(b))


def mismatch_report(a, b, name):
    a = np.ascontiguousarray(a, np.float32)
    b = np.ascontiguousarray(b, np.float32)
    bad = _canon(a) != _canon(b)
    if not bad.any():
        return ""
    idx = np.argwhere(bad)[:5]
    with np.errstate(invalid="ignore"):
        d = np.nanmax(np.abs(a.astype(np.float64) - b.astype(np.float64)))
    return "%s: %d/%d values differ (max abs %.3g); first at %s: %s vs %s" % (
        name, int(bad.sum()), bad.size, d, idx.tolist(), a[bad][:5], b[bad][:5])
