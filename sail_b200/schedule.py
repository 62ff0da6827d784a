"""Deterministic frame schedule.

The reference seeds its shader RNG with wall-clock seconds and jitters the frustum with
``Math.random()`` (/root/reference/src/core/tracer.js:94-98), so its frames are not reproducible.
Here both are explicit: frame k has ``timeSinceStart = float32((k+1)/60)`` and jitter
``((2u1-1)/W, (2u2-1)/H)`` with (u1,u2) the 2k-th and (2k+1)-th outputs of mulberry32(20180120).
"""
import numpy as np

DEFAULT_SEED = 20180120


def _imul(a, b):
    return (a * b) & 0xFFFFFFFF


class Mulberry32:
    def __init__(self, seed=DEFAULT_SEED):
        self.a = seed & 0xFFFFFFFF

    def next(self):
        self.a = (self.a + 0x6D2B79F5) & 0xFFFFFFFF
        a = self.a
        t = _imul(a ^ (a >> 15), 1 | a)
        t = ((t + _imul(t ^ (t >> 7), 61 | t)) & 0xFFFFFFFF) ^ t
        return ((t ^ (t >> 14)) & 0xFFFFFFFF) / 4294967296.0


class FrameSchedule:
    """Supplies (jitter_x, jitter_y, timeSinceStart) for frame k; random access."""

    def __init__(self, seed=DEFAULT_SEED):
        self.seed = seed
        self._rng = Mulberry32(seed)
        self._u = []

    def _ensure(self, count):
        while len(self._u) < count:
            self._u.append(self._rng.next())

    def jitter_unit(self, k):
        self._ensure(2 * k + 2)
        return self._u[2 * k] * 2 - 1, self._u[2 * k + 1] * 2 - 1

    def time(self, k):
        return float(np.float32((k + 1) / 60.0))

    def jitter_units(self, k0, count):
        """(count, 2) array of jitter_unit(k0) .. jitter_unit(k0 + count - 1)."""
        self._ensure(2 * (k0 + count))
        return np.asarray(self._u[2 * k0:2 * (k0 + count)], dtype=np.float64).reshape(count, 2) * 2 - 1

    def times(self, k0, count):
        return ((np.arange(k0, k0 + count, dtype=np.float64) + 1) / 60.0).astype(np.float32)


def frame_matrices(mvp, jitter_units, width, height):
    """tracer.js:94-96 for a batch of frames: inverse(Translation((jx, jy, 0)) * MVP) in double precision, each
    returned as the column-major float32[16] that ``gl.uniformMatrix4fv`` receives (matrix.js:612-622).
    ``jitter_units``: (n, 2) values in [-1, 1).  Translation * MVP only adds jx (jy) times the last row of MVP to
    its first (second) row, which is written out so that one frame and a batch give the same bits."""
    mvp = np.asarray(mvp, dtype=np.float64)
    j = np.asarray(jitter_units, dtype=np.float64).reshape(-1, 2)
    m = np.broadcast_to(mvp, (j.shape[0], 4, 4)).copy()
    m[:, 0, :] += (j[:, 0] / width)[:, None] * mvp[3, :]
    m[:, 1, :] += (j[:, 1] / height)[:, None] * mvp[3, :]
    inv = np.linalg.inv(m)
    return np.ascontiguousarray(inv.transpose(0, 2, 1).reshape(-1, 16).astype(np.float32))


def frame_matrix(mvp, jitter_unit, width, height):
    """One frame of ``frame_matrices``."""
    return frame_matrices(mvp, [jitter_unit], width, height)[0]
