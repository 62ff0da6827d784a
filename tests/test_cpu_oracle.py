"""The oracle against (1) the committed golden fixtures and (2) known-answer tests with analytic results.

The reference has no tests or golden vectors and cannot be run on this box, so parity is
UNPINNED against the reference itself; these tests pin the oracle's behaviour to independent analytic
answers and to a committed fixture (tests/golden/oracle_v1.npz, made by tests/golden/make_golden.py)."""
import math
import os

import numpy as np
import pytest

import sail_b200 as Sail
from conftest import ROOT, bits_equal, frame_uniforms, scene_tables
from sail_b200 import schedule, wire

GOLD = np.load(os.path.join(ROOT, "tests", "golden", "oracle_v1.npz"))
SIZES = {1: (32, 32), 2: (48, 27), 3: (48, 27), 5: (48, 27)}


@pytest.mark.parametrize("cfg", [1, 2, 3, 5])
def test_oracle_matches_golden(oracle, cfg):
    W, H = SIZES[cfg]
    scene, o, t, l, bounces = scene_tables(cfg, W, H)
    assert np.array_equal(o, GOLD["c%d_objects" % cfg]) and np.array_equal(t, GOLD["c%d_texparams" % cfg])
    assert np.array_equal(l, GOLD["c%d_lights" % cfg])
    oracle.set_scene(o, t, l)
    acc = np.zeros((H, W, 3), np.float32)
    for k in range(2):
        m, eye, w, tm = frame_uniforms(scene, k, W, H)
        r = oracle.render_frame(W, H, m, eye, w, tm, acc, max_bounces=bounces, aovs=True)
    assert np.array_equal(r["hit_id"], GOLD["c%d_hit_id" % cfg])
    assert bits_equal(r["depth"], GOLD["c%d_depth" % cfg]) and bits_equal(r["normal"], GOLD["c%d_normal" % cfg])
    assert bits_equal(acc, GOLD["c%d_color" % cfg])


def test_detmath_golden_and_accuracy(oracle):
    xs, us = GOLD["dm_x"], GOLD["dm_u"]
    for name in ("sin", "cos", "tan", "atan", "exp"):
        assert bits_equal(oracle.detmath(name, xs), GOLD["dm_" + name]), name
    assert bits_equal(oracle.detmath("acos", us), GOLD["dm_acos"]) and bits_equal(oracle.detmath("asin", us), GOLD["dm_asin"])
    rng = np.random.default_rng(1)
    x = rng.uniform(-3e5, 3e5, 200000).astype(np.float32)
    assert np.max(np.abs(oracle.detmath("sin", x) - np.sin(x.astype(np.float64)))) < 4e-7
    assert np.max(np.abs(oracle.detmath("cos", x) - np.cos(x.astype(np.float64)))) < 4e-7
    u = rng.uniform(-1, 1, 100000).astype(np.float32)
    assert np.max(np.abs(oracle.detmath("acos", u) - np.arccos(u.astype(np.float64)))) < 5e-7
    a = rng.normal(size=100000).astype(np.float32)
    b = rng.normal(size=100000).astype(np.float32)
    assert np.max(np.abs(oracle.detmath("atan2", a, b) - np.arctan2(a.astype(np.float64), b.astype(np.float64)))) < 5e-7
    e = rng.uniform(-20, 20, 100000).astype(np.float32)
    assert np.max(np.abs(oracle.detmath("exp", e) / np.exp(e.astype(np.float64)) - 1)) < 3e-7
    p = rng.uniform(1e-3, 1e3, 100000).astype(np.float32)
    assert np.max(np.abs(oracle.detmath("log", p) - np.log(p.astype(np.float64)))) < 3e-7


def test_rng_matches_formula(oracle):
    """random.glsl:5-10 in float64 vs the FP32 oracle: the hash amplifies sin() error by 4.4e4, so only a
    loose agreement is expected -- but values must be in [0,1) and equal the golden bits."""
    pts = [(0, 0, 1.0), (17, 5, 2.0166667), (511, 511, 5.5), (1919, 1079, 3.25)]
    got = np.stack([oracle.random2(*p) for p in pts])
    assert bits_equal(got, GOLD["rng"])
    assert np.all(got >= 0) and np.all(got < 1)
    f32 = np.float32
    for (x, y, s), g in zip(pts, got):
        # FP32 dot product step by step (the argument reaches 1e5, where 1 ulp is 0.008 rad), float64 sin
        s = f32(s)
        ax, ay, az = f32(x + 0.5) + s, f32(y + 0.5) + s, f32(0.5) + s
        arg = f32(f32(ax * f32(12.9898)) + f32(ay * f32(78.233))) + f32(az * f32(151.7182))
        v = f32(f32(math.sin(float(arg))) * f32(43758.5453)) + s
        want = float(v) - math.floor(float(v))
        d = abs(want - float(g[0]))
        assert min(d, 1 - d) < 0.02  # sin error 1e-7 x 43758.5


def _one_object(obj, material=None, texture=None):
    scene = Sail.Scene()
    scene.add(obj)
    return wire.serialize_scene(scene)


def test_intersection_known_answers(oracle):
    m, w = Sail.Matte(0.7), Sail.Color.WHITE
    # sphere radius 1 at origin, ray from z=-5 along +z: t = 4, normal (0,0,-1)
    oracle.set_scene(*_one_object(Sail.Sphere([0, 0, 0], 1.0, m, w)))
    ids, t, n, mat = oracle.intersect_rays(np.float32([[0, 0, -5, 0, 0, 1], [0, 0, -5, 0, 0, 2], [0, 0, 0, 1, 0, 0], [0, 2, -5, 0, 0, 1], [0, 0, 0, 0, 1, 0]]))
    assert list(ids) == [0, 0, 0, -1, 0]
    assert t[0] == pytest.approx(4.0, abs=1e-6) and t[1] == pytest.approx(2.0, abs=1e-6)  # t scales with |dir| (Q6)
    assert np.allclose(n[0], [0, 0, -1], atol=1e-6)
    assert t[2] == pytest.approx(1.0, abs=1e-6) and np.allclose(n[2], [-1, 0, 0], atol=1e-6)  # from inside: normal flipped
    assert t[4] == pytest.approx(1.0, abs=1e-6) and np.all(np.isnan(n[4]))  # world +y is the parametric pole: NaN normal (Q18)
    assert t[3] == np.float32(1e5) and mat[0] == 1
    # cube [0,1]^3 from outside (near face) and inside (far face); Cornell box always the far face
    oracle.set_scene(*_one_object(Sail.Cube([0, 0, 0], [1, 1, 1], m, w)))
    ids, t, n, _ = oracle.intersect_rays(np.float32([[0.5, 0.5, -2, 0, 0, 1], [0.5, 0.5, 0.5, 1, 0, 0]]))
    assert t[0] == pytest.approx(2.0, abs=1e-6) and np.allclose(n[0], [0, 0, -1])
    assert t[1] == pytest.approx(0.5, abs=1e-6) and np.allclose(n[1], [-1, 0, 0])
    oracle.set_scene(*_one_object(Sail.Cornellbox([0, 0, 0], [1, 1, 1])))
    ids, t, n, _ = oracle.intersect_rays(np.float32([[0.5, 0.5, -2, 0, 0, 1]]))
    assert t[0] == pytest.approx(3.0, abs=1e-6) and np.allclose(n[0], [0, 0, -1])
    d = oracle.intersect_detail(np.float32([0.5, 0.5, 0.5, -1, 0, 0]))
    assert np.allclose(d["sc"], [0.25, 0.75, 0.25]) and d["into"]  # x- wall is GREEN (cornellbox.glsl:23-37)
    # rectangle in the plane y = 1 spanning x,z in [0,2]; disk of radius 1 (inner 0.5) at the origin in y = 0
    oracle.set_scene(*_one_object(Sail.Rectangle([0, 1, 0], [2, 1, 2], m, w)))
    ids, t, _, _ = oracle.intersect_rays(np.float32([[1, 3, 1, 0, -1, 0], [3, 3, 1, 0, -1, 0]]))
    assert list(ids) == [0, -1] and t[0] == pytest.approx(2.0, abs=1e-6)
    oracle.set_scene(*_one_object(Sail.Disk([0, 0, 0], 1.0, 0.5, m, w)))
    ids, _, _, _ = oracle.intersect_rays(np.float32([[0.75, 1, 0, 0, -1, 0], [0.2, 1, 0, 0, -1, 0], [1.2, 1, 0, 0, -1, 0]]))
    assert list(ids) == [0, -1, -1]
    # cylinder radius 0.5 height 2 on the y axis; cone with apex at y = h
    oracle.set_scene(*_one_object(Sail.Cylinder([0, 0, 0], 2.0, 0.5, m, w)))
    ids, t, n, _ = oracle.intersect_rays(np.float32([[-3, 1, 0, 1, 0, 0], [-3, 2.5, 0, 1, 0, 0]]))
    assert list(ids) == [0, -1] and t[0] == pytest.approx(2.5, abs=1e-5) and np.allclose(n[0], [-1, 0, 0], atol=1e-5)
    oracle.set_scene(*_one_object(Sail.Cone([0, 0, 0], 2.0, 1.0, m, w)))
    ids, t, _, _ = oracle.intersect_rays(np.float32([[-3, 1, 0, 1, 0, 0]]))
    assert ids[0] == 0 and t[0] == pytest.approx(2.5, abs=1e-5)  # radius at half height is 0.5


def test_nearest_hit_tie_break_and_emission_mask(oracle):
    """shader.shape.js:36,47-49: equal distances keep the lower index; emission only on the facing side."""
    m, w = Sail.Matte(0.7), Sail.Color.WHITE
    scene = Sail.Scene()
    scene.add(Sail.Cube([0, 0, 0], [1, 1, 1], m, w, [5, 5, 5]))
    scene.add(Sail.Cube([0, 0, 0], [1, 1, 1], m, w))
    oracle.set_scene(*wire.serialize_scene(scene))
    assert oracle.intersect_rays(np.float32([[0.5, 0.5, -2, 0, 0, 1]]))[0][0] == 0
    out = oracle.intersect_detail(np.float32([0.5, 0.5, -2, 0, 0, 1]))
    assert np.allclose(out["emission"], [5, 5, 5])
    inside = oracle.intersect_detail(np.float32([0.5, 0.5, 0.5, 0, 0, 1]))
    assert np.allclose(inside["emission"], [0, 0, 0])  # back side of the emitter: masked


def test_progressive_mean_is_arithmetic_mean(oracle):
    """fstrace.glsl:14 with textureWeight = k/(k+1) is the running arithmetic mean (what makes SUM sharding legal)."""
    W, H = 24, 24
    scene, o, t, l, bounces = scene_tables(1, W, H)
    oracle.set_scene(o, t, l)
    acc = np.zeros((H, W, 3), np.float32)
    samples = []
    for k in range(6):
        m, eye, w, tm = frame_uniforms(scene, k, W, H)
        oracle.render_frame(W, H, m, eye, w, tm, acc, max_bounces=bounces)
        samples.append(oracle.trace_dump(W, H, m, eye, tm, max_bounces=bounces)["radiance"].astype(np.float64))
    assert np.allclose(acc, np.mean(samples, axis=0), atol=2e-6)


def test_present_window_table_quirk(oracle):
    """window.glsl:38 reads weights[i*j+j]; a constant image must come back constant for every window filter."""
    from sail_b200.filters import lower_filter
    img = np.full((16, 16, 3), 0.25, np.float32)
    pp = Sail.PluginParams("gaussian")
    pp.addParam("r", "vec2(2.0,2.0)")
    pp.addParam("alpha", "2.0")
    f = lower_filter(pp)
    assert len(f["weights"]) == 16 and f["weights"][0] > f["weights"][15] >= 0
    out = oracle.present(img, img, img, f)
    assert np.allclose(out[..., :3], 0.25, atol=1e-6) and np.all(out[..., 3] == 1)
    box = lower_filter(_params("box", r="vec2(2.0,2.0)"))
    assert box["weights"] == [1.0] * 16
    with pytest.raises(ValueError):
        lower_filter(Sail.PluginParams("sinc"))  # missing params: the reference's shader would not compile


def _params(name, **kw):
    pp = Sail.PluginParams(name)
    for k, v in kw.items():
        pp.addParam(k, v)
    return pp


def test_fresnel_known_answers(oracle):
    """fresnel.glsl:31-70 against float64 closed forms."""
    assert oracle.unit("frDielectric", 1.0, 1.0, 1.5) == pytest.approx(0.04, rel=1e-6)           # ((n-1)/(n+1))^2
    assert oracle.unit("frDielectric", 1.0, 1.0, 1.0) == pytest.approx(0.0, abs=1e-12)
    assert oracle.unit("frDielectric", 0.2, 1.5, 1.0) == 1.0                                     # total internal reflection
    c = math.cos(math.radians(56.30993247))                                                      # Brewster angle of n = 1.5
    ct = math.sqrt(1 - (math.sin(math.radians(56.30993247)) / 1.5) ** 2)
    rs = (c - 1.5 * ct) / (c + 1.5 * ct)
    assert oracle.unit("frDielectric", c, 1.0, 1.5) == pytest.approx(0.5 * rs * rs, rel=1e-4)    # r_parallel vanishes
    # conductor at normal incidence: ((n-1)^2 + k^2) / ((n+1)^2 + k^2), per channel
    eta, k = np.array([0.2, 0.92, 1.1]), np.array([3.9, 2.45, 2.14])
    want = ((eta - 1) ** 2 + k ** 2) / ((eta + 1) ** 2 + k ** 2)
    got = oracle.unit("frConductor", 1.0, [1, 1, 1], eta, k)
    assert np.allclose(got, want, rtol=2e-6)
    # grazing incidence reflects everything
    assert np.allclose(oracle.unit("frConductor", 1e-4, [1, 1, 1], eta, k), 1.0, atol=2e-3)


def test_trowbridge_reitz_known_answers(oracle):
    """microfacet.glsl:41-71: D(n) = 1/(pi ax ay); D integrates to 1 against cos; sample_wh uses u.x only (Q11)."""
    assert oracle.unit("trowbridgeReitz_d", 0.3, 0.3, [0, 0, 1]) == pytest.approx(1 / (math.pi * 0.09), rel=1e-6)
    assert oracle.unit("trowbridgeReitz_d", 0.2, 0.5, [0, 0, 1]) == pytest.approx(1 / (math.pi * 0.1), rel=1e-6)
    assert oracle.unit("trowbridgeReitz_d", 0.3, 0.3, [1, 0, 0]) == pytest.approx(0.001)          # tan2Theta = inf sentinel
    # midpoint quadrature of D(wh) cos(theta) over the hemisphere = 1
    ax, ay, n = 0.4, 0.25, 96
    th = (np.arange(n) + 0.5) * (math.pi / 2 / n)
    ph = (np.arange(2 * n) + 0.5) * (math.pi / n)
    total = 0.0
    for t in th:
        for p in ph[::4]:
            wh = [math.sin(t) * math.cos(p), math.sin(t) * math.sin(p), math.cos(t)]
            total += float(oracle.unit("trowbridgeReitz_d", ax, ay, wh)) * math.cos(t) * math.sin(t)
    total *= (math.pi / 2 / n) * (4 * math.pi / n)
    assert total == pytest.approx(1.0, abs=0.02)
    wo = [0.3, 0.1, 0.9]
    a = oracle.unit("trowbridgeReitz_sample_wh", [0.37, 0.11], 0.3, 0.3, wo)
    b = oracle.unit("trowbridgeReitz_sample_wh", [0.37, 0.93], 0.3, 0.3, wo)
    assert np.array_equal(a, b) and a[2] > 0 and np.linalg.norm(a) == pytest.approx(1.0, abs=1e-6)   # u.y unused
    # isotropic: cos(theta) = 1/sqrt(1 + a^2 u/(1-u)), phi = 2 pi u
    ct = 1 / math.sqrt(1 + 0.09 * 0.37 / 0.63)
    assert a[2] == pytest.approx(ct, rel=1e-6) and math.atan2(a[1], a[0]) == pytest.approx(2 * math.pi * 0.37, abs=1e-5)
    assert oracle.unit("trowbridgeReitz_pdf", 0.3, 0.3, wo, a) == pytest.approx(
        float(oracle.unit("trowbridgeReitz_d", 0.3, 0.3, a)) * abs(float(a[2])), rel=1e-6)


def test_samplers_and_quadratic_known_answers(oracle):
    """sampler.glsl and utility.glsl:37-51."""
    for u in ([0.25, 0.6], [0.9, 0.05], [0.0, 0.0]):
        w = oracle.unit("cosineSampleHemisphere", u)
        assert w[2] == pytest.approx(math.sqrt(1 - u[0]), abs=1e-6) and np.linalg.norm(w) == pytest.approx(1.0, abs=1e-6)
        assert math.atan2(w[1], w[0]) % (2 * math.pi) == pytest.approx((2 * math.pi * u[1]) % (2 * math.pi), abs=1e-5)
        s = oracle.unit("uniformSampleSphere", u)
        assert s[2] == pytest.approx(1 - 2 * u[0], abs=1e-6) and np.linalg.norm(s) == pytest.approx(1.0, abs=1e-6)
    assert np.array_equal(oracle.unit("concentricSampleDisk", [0.5, 0.5]), [0, 0])
    d = oracle.unit("concentricSampleDisk", [1.0, 0.5])
    assert d[0] == pytest.approx(1.0) and d[1] == pytest.approx(0.0, abs=1e-7)
    ok, t0, t1 = oracle.unit("quadratic", 1.0, -3.0, 2.0)
    assert ok == 1 and (t0, t1) == (1.0, 2.0)
    assert oracle.unit("quadratic", 1.0, 0.0, 1.0)[0] == 0                                        # negative discriminant
    ok, t0, t1 = oracle.unit("quadratic", 1.0, -1e4, 1.0)                                         # stable small root: C/q
    assert ok == 1 and t0 == pytest.approx(1e-4, rel=1e-6) and t1 == pytest.approx(1e4, rel=1e-6)


def test_textures_known_answers(oracle):
    """texture/*.glsl through getSurfaceColor on a hand-built texParams table."""
    import sail_b200 as Sail
    tex = [Sail.UniformColor([0.1, 0.2, 0.3]), Sail.Checkerboard(0.25, 0.05), Sail.Checkerboard2([1, 0, 0], [0, 0, 1], 0.5),
           Sail.Mix([1, 0, 0], [0, 0, 1], 0.25), Sail.Scale([0.5, 0.5, 1.0], [0.5, 1.0, 0.2]), Sail.UV()]
    rows = np.array([t.gen() for t in tex], np.float32)
    oracle.set_scene(np.zeros((0, 18), np.float32), rows, np.zeros((0, 18), np.float32))
    hit = [0, 0, 0]
    assert np.allclose(oracle.unit("getSurfaceColor", hit, [0.3, 0.7], 0), [0.1, 0.2, 0.3])
    # checkerboard.glsl:6-20: WHITE inside a cell, GREY (0.5) on a grid line of width lineWidth every `size`
    assert np.allclose(oracle.unit("getSurfaceColor", hit, [0.125, 0.125], 1), 1.0)
    assert np.allclose(oracle.unit("getSurfaceColor", hit, [0.251, 0.125], 1), 0.5)
    # checkerboard2.glsl:7-16: parity of floor(u/s) + floor(v/s)
    c00, c10 = oracle.unit("getSurfaceColor", hit, [0.2, 0.2], 2), oracle.unit("getSurfaceColor", hit, [0.7, 0.2], 2)
    assert not np.array_equal(c00, c10) and {tuple(c00), tuple(c10)} == {(1.0, 0.0, 0.0), (0.0, 0.0, 1.0)}
    assert np.array_equal(oracle.unit("getSurfaceColor", hit, [0.7, 0.7], 2), c00)
    # mixf.glsl:7-13 / scale.glsl:6-11 / uvf.glsl:1-3
    m = oracle.unit("getSurfaceColor", hit, [0.3, 0.3], 3)
    assert np.allclose(m, [0.75, 0.0, 0.25], atol=1e-6) or np.allclose(m, [0.25, 0.0, 0.75], atol=1e-6)
    assert np.allclose(oracle.unit("getSurfaceColor", hit, [0.3, 0.3], 4), [0.25, 0.5, 0.2])
    assert np.allclose(oracle.unit("getSurfaceColor", hit, [1.25, -0.25], 5), [0.25, 0.75, 0.0])


def test_light_arms_known_answers(oracle):
    """point.glsl / spot.glsl / area.glsl on an empty scene (nothing to occlude): closed forms of SURVEY 8(a)."""
    import sail_b200 as Sail
    matte = Sail.Matte(0.7)
    scene = Sail.Scene()
    scene.add(Sail.PointLight([0, 4, 0], [0.6, 0.5, 0.4]))
    scene.add(Sail.SpotLight([0, 5, 0], 35, 10, [30, 30, 30]))
    scene.add(Sail.Camera([0, 1, -5], [0, 1, 0]))
    from sail_b200.wire import serialize_scene
    o, t, l = serialize_scene(scene)
    oracle.set_scene(o, t, l)
    hit, n = [0.0, 0.0, 0.0], [0.0, 1.0, 0.0]
    # point light: E * max(0, w.n), no falloff; the sampled point lies within 0.1 of `from`
    p = oracle.unit("light_arm", 10, 20, 0.5, hit, n, 0)
    assert np.all(p <= np.float32([0.6, 0.5, 0.4]) * 1.0000001) and np.allclose(p, [0.6, 0.5, 0.4], rtol=2e-3)
    assert np.array_equal(oracle.unit("light_arm", 10, 20, 0.5, hit, [0, -1, 0], 0), [0, 0, 0])   # facing away
    # spot light straight above: falloff 1 inside the inner cone, E / d^2
    s = oracle.unit("light_arm", 10, 20, 0.5, hit, n, 1)
    assert np.allclose(s, 30.0 / 25.0, rtol=1e-6)
    # outside the cone (35 deg total): tan(40 deg) * 5 away on the floor
    far = [5 * math.tan(math.radians(40)), 0.0, 0.0]
    assert np.array_equal(oracle.unit("light_arm", 10, 20, 0.5, far, n, 1), [0, 0, 0])
    # in the falloff band (between 25 and 35 deg): delta^4 smooth step, strictly between 0 and the full value
    mid = [5 * math.tan(math.radians(30)), 0.0, 0.0]
    v = oracle.unit("light_arm", 10, 20, 0.5, mid, n, 1)
    d2 = 25 + mid[0] ** 2
    cosw = 5 / math.sqrt(d2)
    delta = (cosw - math.cos(math.radians(35))) / (math.cos(math.radians(25)) - math.cos(math.radians(35)))
    assert np.allclose(v, 30.0 * delta ** 4 * cosw / d2, rtol=5e-5)


def test_oracle_matches_golden_v2_bounce_dump(oracle):
    """tests/golden/oracle_v2.npz: per-bounce rays / hits of frame 0 of config 2 (the converged images of the same file
    take ~10 s of CPU and are regenerated by tests/golden/make_golden_v2.py; the GPU suite checks them)."""
    gold = np.load(os.path.join(ROOT, "tests", "golden", "oracle_v2.npz"))
    W, H = 48, 27
    scene, o, t, l, bounces = scene_tables(2, W, H)
    oracle.set_scene(o, t, l)
    m, eye, w, tm = frame_uniforms(scene, 0, W, H)
    d = oracle.trace_dump(W, H, m, eye, tm, max_bounces=bounces)
    assert np.array_equal(d["hit_id"], gold["c2_bounce_hit_id"])
    for k in ("t", "rays", "radiance"):
        assert bits_equal(d[k], gold["c2_bounce_" + k]), k
    assert gold["c1_converged"].shape == (64, 64, 3) and np.isfinite(gold["c1_converged"]).all()
    assert 0.05 < float(gold["c1_converged"].mean()) < 1.0


def test_oracle_reproduces_golden_v3(oracle):
    """The oracle against its committed round-2 fixture (tests/golden/oracle_v3.npz): plugin-matrix scenes (Oren-Nayar, rough
    glass, Bilerp / Scale, sphere / disk / stub area lights, reverseNormal, single-row tables) and the opt-in Russian roulette.
    Guards the oracle -- and the shared detmath.h -- against silent changes; the full-size digests are checked on the GPU box."""
    import os
    import matrix_scenes
    from conftest import ROOT, frame_uniforms, scene_tables
    from sail_b200 import wire
    gold = np.load(os.path.join(ROOT, "tests", "golden", "oracle_v3.npz"))
    W, H = 48, 32
    for name, fn in sorted(matrix_scenes.MATRIX.items()):
        scene = fn(W, H)
        oracle.set_scene(*wire.serialize_scene(scene))
        acc = np.zeros((H, W, 3), np.float32)
        for k in range(3):
            m, eye, w, tm = frame_uniforms(scene, k, W, H)
            r = oracle.render_frame(W, H, m, eye, w, tm, acc, max_bounces=5, aovs=True)
        assert np.array_equal(acc.view(np.uint32), gold["m_%s_color" % name].view(np.uint32)), name
        assert np.array_equal(r["hit_id"], gold["m_%s_hit_id" % name].astype(np.int32)), name
    scene, o, t, l, bounces = scene_tables(2, W, H)
    oracle.set_scene(o, t, l)
    oracle.set_russian_roulette(2)
    try:
        acc = np.zeros((H, W, 3), np.float32)
        for k in range(4):
            m, eye, w, tm = frame_uniforms(scene, k, W, H)
            oracle.render_frame(W, H, m, eye, w, tm, acc, max_bounces=bounces)
    finally:
        oracle.set_russian_roulette(0)
    assert np.array_equal(acc.view(np.uint32), gold["rr2_c2_color"].view(np.uint32))
