"""Wire-format conformance: the rows every gen() emits, checked against row layouts
written out by hand from /root/reference/src/scene/{geometry,material,texture,light}.js and appendix A."""
import math

import numpy as np
import pytest

import sail_b200 as Sail
from sail_b200 import schedule, wire


def row(vals, n):
    return np.asarray(list(vals) + [0.0] * (n - len(vals)), np.float32)


def test_object_rows():
    m, t = Sail.Matte(0.7), Sail.UniformColor([1, 1, 1])
    # geometry.js:82-90: tail = reverseNormal, matRow, texRow, emission; pad to 18
    assert np.array_equal(np.float32(Sail.Cube([1, 2, 3], [4, 5, 6], m, t, [7, 8, 9], True).gen(4)),
                          row([1, 1, 2, 3, 4, 5, 6, 1, 4, 5, 7, 8, 9], 18))
    assert np.array_equal(np.float32(Sail.Sphere([1, 2, 3], 0.5, m, t).gen(2)), row([2, 1, 2, 3, 0.5, 0, 2, 3, 0, 0, 0], 18))
    assert np.array_equal(np.float32(Sail.Rectangle([0, 1, 2], [3, 1, 5], m, t).gen(0)), row([3, 0, 1, 2, 3, 1, 5, 0, 0, 1], 18))
    assert np.array_equal(np.float32(Sail.Cone([1, 0, 2], 1.6, 0.7, m, t).gen(6)), row([4, 1, 0, 2, 1.6, 0.7, 0, 6, 7], 18))
    assert np.array_equal(np.float32(Sail.Cylinder([1, 0, 2], 1.2, 0.5, m, t).gen(6)), row([5, 1, 0, 2, 1.2, 0.5, 0, 6, 7], 18))
    assert np.array_equal(np.float32(Sail.Disk([1, 0, 2], 0.8, 0.3, m, t).gen(8)), row([6, 1, 0, 2, 0.8, 0.3, 0, 8, 9], 18))
    assert np.array_equal(np.float32(Sail.Paraboloid([1, 0, 2], 0, 1.0, 0.6, m, t).gen(2)), row([8, 1, 0, 2, 0, 1.0, 0.6, 0, 2, 3], 18))
    # Cornellbox: default min/max geometry.js:611; generic tail (which cornellbox.glsl:17 then mis-reads, Q3)
    assert np.array_equal(np.float32(Sail.Cornellbox().gen(2)), row([9, 0, 0, -5, 5.560, 5.488, 5.592, 0, 2, 3], 18))


def test_hyperboloid_fit():
    """geometry.js:467-486: ah, ch fitted on the host; SURVEY.md 8(d) config 3 gives ah=6.25, ch=1.5625."""
    h = Sail.Hyperboloid([2.78, 1.0, 3.0], [0.4, -0.3, -0.6], [0.4, 0.3, 0.6], Sail.Metal(0, 0.05, 0.05), Sail.Color.WHITE)
    r = h.gen(0)
    assert r[0] == 7 and r[1:4] == [2.78, 1.0, 3.0]
    assert r[10] == pytest.approx(6.25, rel=1e-12) and r[11] == pytest.approx(1.5625, rel=1e-12)
    assert r[12:15] == [0.0, 0.0, 1.0]
    with pytest.raises(Exception):
        Sail.Hyperboloid([0, 0, 0], [0, 0, 0], [0, 0, 0], Sail.Matte(), Sail.Color.WHITE)


def test_material_rows():
    assert np.array_equal(np.float32(Sail.Matte(0.7).gen()), row([1, 0.7, 0, 0, 0], 16))
    assert Sail.Matte(-1).gen()[1] == 1  # material.js:36
    s = 20 * math.pi / 180
    on = Sail.Matte(0.5, 20).gen()
    assert on[2] == 20 and on[3] == pytest.approx(1 - s * s / (2 * (s * s + 0.33))) and on[4] == pytest.approx(0.45 * s * s / (s * s + 0.09))
    assert np.array_equal(np.float32(Sail.Mirror(1.0).gen()), row([2, 1.0], 16))
    assert Sail.Mirror(0).gen()[1] == 0.5  # material.js:65
    # material.js:84-88: u/v roughness 0 -> roughness; default eta / k
    assert np.allclose(Sail.Metal(0, 0.01, 0.1).gen()[:9], [3, 0.01, 0.1, 9.5308, 6.6358, 4.4751, 13.0282, 8.1126, 5.5028])
    assert Sail.Metal(0.2).gen()[1:3] == [0.2, 0.2]
    assert np.array_equal(np.float32(Sail.Glass(1, 1, 1.5).gen()), row([4, 1, 1, 1.5, 0, 0], 16))


def test_texture_rows():
    assert np.array_equal(np.float32(Sail.UniformColor([0.1, 0.2, 0.3]).gen()), row([0, 0.1, 0.2, 0.3], 16))
    assert np.array_equal(np.float32(Sail.Checkerboard(0.1, 0.01).gen()), row([5, 0.1, 0.01], 16))
    assert np.array_equal(np.float32(Sail.Checkerboard2([1, 1, 1], [0, 0, 0], 0.1).gen()), row([7, 1, 1, 1, 0, 0, 0, 0.1], 16))
    assert np.array_equal(np.float32(Sail.Bilerp([1, 0, 0], [0, 1, 0], [0, 0, 1], [1, 1, 1]).gen()), row([8, 1, 0, 0, 0, 1, 0, 0, 0, 1, 1, 1, 1], 16))
    assert np.array_equal(np.float32(Sail.Mix([0.9, 0.2, 0.2], [0.2, 0.2, 0.9], 0.35).gen()), row([9, 0.9, 0.2, 0.2, 0.2, 0.2, 0.9, 0.35], 16))
    assert np.array_equal(np.float32(Sail.Scale([1, 2, 3], [4, 5, 6]).gen()), row([10, 1, 2, 3, 4, 5, 6], 16))
    assert np.array_equal(np.float32(Sail.UV().gen()), row([11], 16))


def test_light_rows_and_geometry_injection():
    scene = Sail.Scene()
    scene.add(Sail.Cornellbox())
    rect = Sail.Rectangle([2.13, 5.48, 2.27], [3.43, 5.48, 3.32], Sail.Matte(0.7), Sail.Color.BLACK)
    area = Sail.AreaLight(rect, [4, 4, 4])
    with pytest.raises(RuntimeError):
        area.gen()  # light.js:52: not attached to a scene yet
    scene.add(area)  # scene.js:57-61: the geometry is pushed into scene.objects, index = its position
    scene.add(Sail.PointLight([0.8, 4.6, 0.8], [0.6, 0.6, 0.6]))
    scene.add(Sail.SpotLight([4.6, 5.2, 2.0], 35, 10, [30, 30, 30]))
    o, t, l = wire.serialize_scene(scene)
    assert o.shape == (2, 18) and t.shape == (4, 16) and l.shape == (3, 18)
    assert o[1, 0] == 3 and list(o[1, 10:13]) == [4, 4, 4]          # emission copied onto the geometry (light.js:31-35)
    assert list(o[:, 8]) == [0, 2] and list(o[:, 9]) == [1, 3]       # matRow = 2i, texRow = 2i+1 (tracer.js:46-49)
    assert np.array_equal(l[0], row([0, 1, 4, 4, 4], 18))
    assert np.array_equal(l[1], row([1, 0.8, 4.6, 0.8, 0.6, 0.6, 0.6], 18))
    assert np.allclose(l[2, :9], [2, math.cos(35 / 180 * math.pi), math.cos(25 / 180 * math.pi), 4.6, 5.2, 2.0, 30, 30, 30])
    # updateObjects re-uses the stored texparamID (tracer.js:25-31)
    assert np.array_equal(wire.serialize_objects(scene), o)


def test_plugin_discovery_and_params():
    from sail_b200.scene import plugin_masks
    scene, *_ = __import__("sail_b200.configs", fromlist=["x"]).config3(64, 64)
    cfg = scene.tracerConfig()
    assert [p.name for p in cfg["shape"]] == ["cornellbox", "cone", "cylinder", "hyperboloid", "paraboloid", "disk", "rectangle"]
    assert [p.name for p in cfg["light"]] == ["area", "point", "spot"]
    m = plugin_masks(cfg)
    assert m["material_mask"] == (1 << 1) | (1 << 3) | (1 << 4) and m["light_mask"] == 0b111
    scene.filter = "no-such-filter"  # scene.js:24-31: unknown names are ignored
    assert scene.filter.name == "tonemapping"
    pp = Sail.PluginParams("gaussian")
    pp.addParam("r", "vec2(2.25,3.0)")
    assert pp.getParam("r") == [2.2, 3.0]  # generator.js:14-20 keeps one fractional digit (Q10)


def test_frame_schedule_and_matrix():
    s = schedule.FrameSchedule()
    u = [s.jitter_unit(k) for k in range(4)]
    assert all(-1 <= a < 1 and -1 <= b < 1 for a, b in u) and len(set(u)) == 4
    assert u == [schedule.FrameSchedule().jitter_unit(k) for k in range(4)]  # random access is deterministic
    assert s.time(0) == float(np.float32(1 / 60.0))
    # mulberry32(20180120): first outputs pinned (computed from the definition in SURVEY.md 8(d))
    r = schedule.Mulberry32(20180120)
    a = (20180120 + 0x6D2B79F5) & 0xFFFFFFFF
    t = ((a ^ (a >> 15)) * (1 | a)) & 0xFFFFFFFF
    t = ((t + (((t ^ (t >> 7)) * (61 | t)) & 0xFFFFFFFF)) & 0xFFFFFFFF) ^ t
    assert r.next() == ((t ^ (t >> 14)) & 0xFFFFFFFF) / 4294967296.0
    cam = Sail.Camera([2.78, 2.73, -6], [2.78, 2.73, 2.79])
    mvp = cam.projection @ cam.modelview
    m = schedule.frame_matrix(mvp, (0.0, 0.0), 512, 512).reshape(4, 4).T  # column-major -> row-major
    assert np.allclose(m @ mvp, np.eye(4), atol=1e-5)
    # camera.js:38-41: X axis negated (mirror image, Q14): a standard lookAt down +z gives x = (-1,0,0)
    assert cam.modelview[0, 0] == pytest.approx(1.0)


def test_batched_frames_equal_single_frames():
    """Tracer.frames (one numpy pass, used by Renderer.trace) must give the very bytes of Tracer.frame per sample."""
    from sail_b200 import capi
    from sail_b200.renderer import Tracer
    from sail_b200.schedule import FrameSchedule

    class Size:
        width, height = 320, 200

    from conftest import scene_tables
    scene, *_ = scene_tables(2, 320, 200)
    tr = Tracer(Size(), FrameSchedule())
    one = capi.frame_array([tr.frame(scene.mat, scene.eye, k) for k in range(7, 7 + 40)])
    many = tr.frames(scene.mat, scene.eye, 7, 40)
    assert bytes(one) == bytes(many)


def test_image_writers_round_trip(tmp_path):
    """PNG/PPM writers of the display output: bottom-up device rows come out top-down."""
    from sail_b200 import image
    rng = np.random.default_rng(3)
    img = rng.integers(0, 256, size=(37, 53, 4), dtype=np.uint8)
    image.write_png(tmp_path / "a.png", img)
    assert np.array_equal(image.read_png(tmp_path / "a.png"), img[::-1, :, :3])
    image.write_png(tmp_path / "b.png", img, alpha=True)
    assert np.array_equal(image.read_png(tmp_path / "b.png"), img[::-1])
    image.write_ppm(tmp_path / "a.ppm", img)
    raw = open(tmp_path / "a.ppm", "rb").read()
    assert raw.startswith(b"P6\n53 37\n255\n") and raw[len(b"P6\n53 37\n255\n"):] == img[::-1, :, :3].tobytes()
    # float display output: clamp + round to nearest like an RGBA8 target
    f = np.array([[[0.0, 0.5, 1.0, 1.0], [-1.0, 2.0, float("nan"), 1.0]]], np.float32)
    image.write_png(tmp_path / "c.png", f)
    assert image.read_png(tmp_path / "c.png").tolist() == [[[0, 128, 255], [0, 255, 0]]]


def test_pickup_ray_and_boundboxes():
    """pickup.js:9-38 and the boundbox() of every shape (geometry.js), host-side arithmetic only."""
    from sail_b200.pickup import MAXVALUE, Ray
    box = {'min': [-1, -1, 4], 'max': [1, 1, 6]}
    r = Ray([0, 0, 0], [0, 0, 2])
    assert r.testBoundBox(box) and r.intersectBoundBox(box) == 2.0      # un-normalised dir: t in units of |dir|
    assert Ray([0, 0, 5], [0, 0, 1]).intersectBoundBox(box) == 1.0       # inside: far face
    assert not Ray([0, 0, 0], [0, 0, -1]).testBoundBox(box)
    assert Ray([0, 5, 0], [0, 0, 1]).intersectBoundBox(box) == MAXVALUE
    m, w = Sail.Matte(0.7), Sail.Color.WHITE
    assert Sail.Cornellbox().boundbox() is False                         # geometry.js:51-53: walls are not pickable
    assert Sail.Sphere([1, 2, 3], 0.5, m, w).boundbox() == {'min': [0.5, 1.5, 2.5], 'max': [1.5, 2.5, 3.5]}
    bb = Sail.Rectangle([0, 1, 0], [2, 1, 3], m, w).boundbox()
    assert bb['min'] == [0, 0.95, 0] and bb['max'] == [2, 1.05, 3]
    assert Sail.Cone([0, 0, 0], 2, 1, m, w).boundbox() == {'min': [-1, 0, -1], 'max': [1, 2, 1]}
    assert Sail.Disk([0, 1, 0], 2, 0.5, m, w).boundbox() == {'min': [-2, 0.95, -2], 'max': [2, 1.05, 2]}
    hb = Sail.Hyperboloid([0, 1, 0], [0.4, -0.3, -0.6], [0.4, 0.3, 0.6], m, w).boundbox()
    assert hb['min'][1] == pytest.approx(0.4) and hb['max'][1] == pytest.approx(1.6) and hb['max'][0] == pytest.approx(0.5)
    pb = Sail.Paraboloid([0, 0, 0], 0, 1, 0.6, m, w).boundbox()
    assert pb == {'min': [-0.6, 0, -0.6], 'max': [0.6, 1, 0.6]}
