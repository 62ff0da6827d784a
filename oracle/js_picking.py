"""TEST INFRASTRUCTURE ONLY -- a second, independent reading of the shapes' geometry.

The only CPU arithmetic the reference contains is the JavaScript picking code: ``Shape.intersect(ray)`` of every shape class
(/root/reference/src/scene/geometry.js) and ``quadratic`` (geometry.js:10-25).  It cannot be executed here (no JS engine),
so it is restated below in float64 numpy, vectorised over rays, line by line.  It is NOT the shader: other epsilons
(``MINVALUE = 1e-4`` / ``MAXVALUE = 1e5``, rollup.config.js:82-85, against the GLSL's ``EPSILON = 1e-5``), double precision,
distance only.  tests/test_cpu_js_picking.py uses it as a loose cross-check of the oracle's intersectors
(|t_js - t_oracle| <= 1e-3 wherever both report a hit beyond 1e-3) -- the one place where parity can lean
on reference-authored arithmetic instead of on the oracle's own output.

Known reference bug, restated as written: ``Rectangle.intersect`` tests ``hit.x`` / ``hit.y`` on a ``Vector`` -- ``x`` is
the alias of the ``multiply`` method (utils/matrix.js:95) and ``y`` is undefined, so every comparison is false and the
bounds never reject (geometry.js:244-245): the JS rectangle is an infinite plane.  The test therefore only checks the
distances where the oracle (the GLSL's bounded rectangle) hits.
"""
import numpy as np

MINVALUE = 1e-4
MAXVALUE = 1e5


def quadratic(A, B, C):
    """geometry.js:10-25.  Returns (t0, t1); (-1, -1) where the discriminant is negative."""
    with np.errstate(all="ignore"):
        discrim = B * B - 4.0 * A * C
        ok = ~(discrim < 0.0)
        root = np.sqrt(np.where(ok, discrim, 0.0))
        q = np.where(B < 0.0, -0.5 * (B - root), -0.5 * (B + root))
        t0 = q / A
        t1 = C / q
        swap = t0 > t1
        t0, t1 = np.where(swap, t1, t0), np.where(swap, t0, t1)
    return np.where(ok, t0, -1.0), np.where(ok, t1, -1.0)


def to_object_space(v, offset):
    """Object3D.toObjectSpace with the default frame n = j, s = -k, t = i (geometry.js:47-58): (v.s, v.t, v.n)."""
    v = v - np.asarray(offset, np.float64)
    return np.stack([-v[:, 2], v[:, 0], v[:, 1]], axis=1)


def _rays(origin, direction, offset):
    o = to_object_space(np.asarray(origin, np.float64), offset)
    d = to_object_space(np.asarray(direction, np.float64), [0.0, 0.0, 0.0])
    return o, d


def cube(o, d, mn, mx):
    """Cube.intersect, geometry.js:110-121."""
    o, d = np.asarray(o, np.float64), np.asarray(d, np.float64)
    with np.errstate(all="ignore"):
        tMin = (np.asarray(mn, np.float64) - o) / d
        tMax = (np.asarray(mx, np.float64) - o) / d
    tNear = np.minimum(tMin, tMax).max(axis=1)
    tFar = np.maximum(tMin, tMax).min(axis=1)
    return np.where((tNear > MINVALUE) & (tNear < tFar), tNear, np.where(tNear < tFar, tFar, MAXVALUE))


def sphere(o, d, c, r):
    """Sphere.intersect, geometry.js:160-174."""
    ro, rd = _rays(o, d, c)
    a = (rd * rd).sum(1)
    b = 2.0 * (ro * rd).sum(1)
    cc = (ro * ro).sum(1) - r * r
    t1, t2 = quadratic(a, b, cc)
    t = np.where(t1 < MINVALUE, t2, t1)
    return np.where(t2 < MINVALUE, MAXVALUE, t)


def rectangle(o, d, mn, mx):
    """Rectangle.intersect, geometry.js:224-248 -- with its bounds test as it actually evaluates (never rejects)."""
    mn, mx = np.asarray(mn, np.float64), np.asarray(mx, np.float64)
    x = np.array([mx[0] - mn[0], 0.0, 0.0])
    y = np.array([0.0, mx[1] - mn[1], mx[2] - mn[2]])
    s = x / np.linalg.norm(x)
    t = y / np.linalg.norm(y)
    n = np.cross(s, t)
    v = np.asarray(o, np.float64) - mn
    ro_z = v @ n
    rd_z = np.asarray(d, np.float64) @ n
    with np.errstate(all="ignore"):
        tt = -ro_z / rd_z
    out = np.where(tt < MINVALUE, MAXVALUE, tt)
    return np.where(rd_z == 0.0, MAXVALUE, out)


def _clipped(t1, t2, ro, rd, zlo, zhi, first_reject):
    """The tail shared by cone / cylinder / hyperboloid / paraboloid (e.g. geometry.js:296-314)."""
    t = np.where(t1 < MINVALUE, t2, t1)
    hz = ro[:, 2] + rd[:, 2] * t
    out = (hz < zlo) | (hz > zhi)
    hz2 = ro[:, 2] + rd[:, 2] * t2
    out2 = (hz2 < zlo) | (hz2 > zhi)
    miss = first_reject | (out & ((t == t2) | out2))
    t = np.where(out, t2, t)
    miss = miss | (t >= MAXVALUE)
    return np.where(miss, MAXVALUE, t)


def cone(o, d, position, height, radius):
    """Cone.intersect, geometry.js:286-315."""
    ro, rd = _rays(o, d, position)
    k = (radius / height) ** 2
    a = rd[:, 0] ** 2 + rd[:, 1] ** 2 - k * rd[:, 2] ** 2
    b = 2.0 * (rd[:, 0] * ro[:, 0] + rd[:, 1] * ro[:, 1] - k * rd[:, 2] * (ro[:, 2] - height))
    c = ro[:, 0] ** 2 + ro[:, 1] ** 2 - k * (ro[:, 2] - height) ** 2
    t1, t2 = quadratic(a, b, c)
    return _clipped(t1, t2, ro, rd, -MINVALUE, height, t2 < -MINVALUE)


def cylinder(o, d, position, height, radius):
    """Cylinder.intersect, geometry.js:352-380."""
    ro, rd = _rays(o, d, position)
    a = rd[:, 0] ** 2 + rd[:, 1] ** 2
    b = 2.0 * (rd[:, 0] * ro[:, 0] + rd[:, 1] * ro[:, 1])
    c = ro[:, 0] ** 2 + ro[:, 1] ** 2 - radius * radius
    t1, t2 = quadratic(a, b, c)
    return _clipped(t1, t2, ro, rd, -MINVALUE, height, t2 < MINVALUE)


def disk(o, d, position, radius, inner):
    """Disk.intersect, geometry.js:417-432."""
    ro, rd = _rays(o, d, position)
    with np.errstate(all="ignore"):
        t = -ro[:, 2] / rd[:, 2]
    hx, hy = ro[:, 0] + rd[:, 0] * t, ro[:, 1] + rd[:, 1] * t
    dist2 = hx * hx + hy * hy
    miss = (rd[:, 2] == 0.0) | (t <= MINVALUE) | (dist2 > radius * radius) | (dist2 < inner * inner) | (t >= MAXVALUE)
    return np.where(miss, MAXVALUE, t)


def hyperboloid(o, d, position, ah, ch, zmin, zmax):
    """Hyperboloid.intersect, geometry.js:496-521 (ah, ch: the constructor's fit, geometry.js:467-486)."""
    ro, rd = _rays(o, d, position)
    a = ah * rd[:, 0] ** 2 + ah * rd[:, 1] ** 2 - ch * rd[:, 2] ** 2
    b = 2.0 * (ah * rd[:, 0] * ro[:, 0] + ah * rd[:, 1] * ro[:, 1] - ch * rd[:, 2] * ro[:, 2])
    c = ah * ro[:, 0] ** 2 + ah * ro[:, 1] ** 2 - ch * ro[:, 2] ** 2 - 1.0
    t1, t2 = quadratic(a, b, c)
    return _clipped(t1, t2, ro, rd, zmin, zmax, t2 < -MINVALUE)


def paraboloid(o, d, position, z0, z1, radius):
    """Paraboloid.intersect, geometry.js:564-591 (zMin / zMax = min / max(z0, z1))."""
    ro, rd = _rays(o, d, position)
    zmin, zmax = min(z0, z1), max(z0, z1)
    k = zmax / (radius * radius)
    a = k * (rd[:, 0] ** 2 + rd[:, 1] ** 2)
    b = 2.0 * k * (rd[:, 0] * ro[:, 0] + rd[:, 1] * ro[:, 1]) - rd[:, 2]
    c = k * (ro[:, 0] ** 2 + ro[:, 1] ** 2) - ro[:, 2]
    t1, t2 = quadratic(a, b, c)
    return _clipped(t1, t2, ro, rd, zmin, zmax, t2 < -MINVALUE)
