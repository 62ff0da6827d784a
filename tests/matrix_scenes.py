"""The "plugin matrix": small scenes that put every arm of the path under test that the BASELINE configs never reach
(VERDICT round 1, items N1/2): Oren-Nayar, rough glass / MicrofacetT, Bilerp, Scale, sphere / disk / stub area lights,
reverseNormal on every shape, uv textures on cube / sphere / hyperboloid / paraboloid, and single-row tables (Q5).

Used by tests/test_gpu_matrix.py (CUDA vs oracle, both kernels, jit + aot) and tests/golden/make_golden_v3.py.
Reference arms: /root/reference/src/shader/material/bsdf.glsl:40-76,198-244, material/glass.glsl:21-33,
texture/bilerp.glsl:1-13, texture/scale.glsl:1-11, shape/sphere.glsl:88-92, shape/disk.glsl:77-83, shape/cube.glsl:50-63,
shape/shader.shape.js:47-49.
"""
import sail_b200 as Sail
from sail_b200 import configs

WHITE = [1, 1, 1]


def _finish(scene, W, H, filter_name="color"):
    scene.add(configs._camera(W, H))
    scene.filter = filter_name
    return scene


def oren_nayar(W, H):
    """Matte(kd, sigma > 0) on walls (Cornell box reads material row 0, Q3), spheres and a cube, under an area light (NEE)."""
    s = Sail.Scene()
    rough = Sail.Matte(0.8, 20)
    s.add(Sail.Cube([0.8, 0.0, 1.0], [2.0, 1.6, 2.2], rough, Sail.UniformColor([0.9, 0.6, 0.3])))  # row 0: the walls' material too
    s.add(Sail.Cornellbox())
    s.add(Sail.Sphere([3.6, 1.0, 2.4], 1.0, Sail.Matte(0.7, 45), Sail.UniformColor([0.3, 0.6, 0.9])))
    s.add(Sail.Sphere([2.6, 0.5, 0.6], 0.5, Sail.Matte(0.7), Sail.Color.WHITE))  # a Lambert one next to them
    s.add(Sail.AreaLight(Sail.Rectangle([2.13, 5.48, 2.27], [3.43, 5.48, 3.32], Sail.Matte(0.7), Sail.Color.BLACK), [3, 3, 3]))
    return _finish(s, W, H, "tonemapping")


def rough_glass(W, H):
    """Glass with uroughness / vroughness > 0: MicrofacetR for u.x < 0.5 (negative remapped u.x, Q11), MicrofacetT above;
    isotropic and anisotropic Trowbridge-Reitz; a smooth glass and a metal next to them."""
    s = Sail.Scene()
    matte = Sail.Matte(0.7)
    s.add(Sail.AreaLight(Sail.Rectangle([2.13, 5.48, 2.27], [3.43, 5.48, 3.32], matte, Sail.Color.BLACK), [2, 2, 2]))
    s.add(Sail.Cornellbox())
    s.add(Sail.Sphere([1.4, 1.2, 2.6], 1.2, Sail.Glass(1, 1, 1.5, 0.2, 0.2), Sail.Color.WHITE))
    s.add(Sail.Sphere([4.0, 1.0, 1.8], 1.0, Sail.Glass(0.9, 0.8, 1.33, 0.05, 0.3), Sail.UniformColor([0.9, 1.0, 0.9])))
    s.add(Sail.Sphere([2.8, 0.5, 0.3], 0.5, Sail.Glass(1, 1, 1.5), Sail.Color.WHITE))
    s.add(Sail.Cube([2.3, 0.0, 3.6], [3.3, 1.0, 4.6], Sail.Metal(0, 0.2, 0.2), Sail.Color.WHITE))
    return _finish(s, W, H, "tonemapping")


def textures(W, H):
    """Bilerp, Scale, and Checkerboard / Checkerboard2 / UV on the shapes whose uv never reached a texture before:
    cube, sphere, hyperboloid, paraboloid (plus cone with Bilerp)."""
    s = Sail.Scene()
    matte = Sail.Matte(0.7)
    s.add(Sail.Cornellbox())
    s.add(Sail.Cube([0.5, 0.0, 2.5], [1.7, 1.2, 3.7], matte, Sail.Checkerboard(0.25, 0.05)))
    s.add(Sail.Sphere([2.8, 1.0, 3.2], 1.0, matte, Sail.Checkerboard2([0.9, 0.9, 0.1], [0.1, 0.1, 0.9], 0.125)))
    s.add(Sail.Hyperboloid([4.4, 1.0, 2.0], [0.5, -0.3, -0.7], [0.5, 0.3, 0.7], matte, Sail.UV()))
    s.add(Sail.Paraboloid([1.2, 0.0, 0.8], 0, 1.0, 0.6, matte, Sail.Bilerp([1, 0, 0], [0, 1, 0], [0, 0, 1], [1, 1, 0])))
    s.add(Sail.Cone([3.0, 0.0, 0.9], 1.2, 0.5, matte, Sail.Bilerp([0.2, 0.2, 0.2], [1, 1, 1], [0.9, 0.1, 0.1], [0.1, 0.9, 0.1])))
    s.add(Sail.Sphere([4.6, 0.5, 4.2], 0.5, matte, Sail.Scale([0.9, 0.5, 0.5], [0.5, 0.9, 1.0])))
    s.add(Sail.Cube([0.4, 2.2, 4.6], [1.4, 3.2, 5.4], matte, Sail.UV()))
    s.add(Sail.AreaLight(Sail.Rectangle([2.13, 5.48, 2.27], [3.43, 5.48, 3.32], matte, Sail.Color.BLACK), [3, 3, 3]))
    return _finish(s, W, H)


def light_shapes(W, H):
    """AreaLight over a Sphere (pdf INVPI/r^2, sphere.glsl:88-92), a Disk (concentric sampling, disk.glsl:77-83) and a
    Cube (stub sampler: BLACK with pdf unwritten, cube.glsl:50-52, Q7).  All three lights are area lights: with the
    category read from row 0 / the last row (Q4) any other kind in the last row would turn lights 1.. into that kind."""
    s = Sail.Scene()
    matte = Sail.Matte(0.7)
    s.add(Sail.Cornellbox())
    s.add(Sail.AreaLight(Sail.Sphere([1.2, 4.2, 2.4], 0.4, matte, Sail.Color.BLACK), [6, 5, 4]))
    s.add(Sail.AreaLight(Sail.Disk([3.8, 5.2, 3.0], 0.7, 0.2, matte, Sail.Color.BLACK), [4, 5, 6]))
    s.add(Sail.AreaLight(Sail.Cube([2.4, 4.9, 0.6], [3.2, 5.2, 1.4], matte, Sail.Color.BLACK), [5, 5, 5]))
    s.add(Sail.Sphere([2.0, 1.0, 3.0], 1.0, matte, Sail.Color.WHITE))
    s.add(Sail.Cube([3.6, 0.0, 1.6], [4.6, 1.4, 2.6], matte, Sail.UniformColor([0.8, 0.7, 0.6])))
    return _finish(s, W, H, "tonemapping")


def reverse_normals(W, H):
    """reverseNormal = true on each shape (shader.shape.js:47-49): emission is seen from the inside only; a reversed
    rectangle / sphere / disk as area-light geometry flips the sampled light normal."""
    s = Sail.Scene()
    matte = Sail.Matte(0.7)
    em = [2, 2, 2]
    s.add(Sail.Cornellbox())
    s.add(Sail.Cube([0.4, 0.0, 3.4], [1.4, 1.0, 4.4], matte, Sail.Color.WHITE, em, True))
    s.add(Sail.Sphere([2.4, 0.8, 3.6], 0.8, matte, Sail.Color.WHITE, em, True))
    s.add(Sail.Cone([4.0, 0.0, 3.6], 1.4, 0.6, matte, Sail.Color.WHITE, em, True))
    s.add(Sail.Cylinder([0.9, 0.0, 1.4], 1.0, 0.5, matte, Sail.Color.WHITE, em, True))
    s.add(Sail.Disk([2.6, 0.02, 1.4], 0.7, 0.2, matte, Sail.Color.WHITE, em, True))
    s.add(Sail.Hyperboloid([4.4, 1.0, 1.4], [0.4, -0.3, -0.6], [0.4, 0.3, 0.6], matte, Sail.Color.WHITE, em, True))
    s.add(Sail.Paraboloid([2.8, 2.0, 4.6], 0, 0.8, 0.5, matte, Sail.Color.WHITE, em, True))
    s.add(Sail.Rectangle([1.0, 2.5, 5.4], [2.0, 3.5, 5.4], matte, Sail.Color.WHITE, em, True))
    s.add(Sail.AreaLight(Sail.Rectangle([2.13, 5.48, 2.27], [3.43, 5.48, 3.32], matte, Sail.Color.BLACK, reverseNormal=True), [3, 3, 3]))
    s.add(Sail.AreaLight(Sail.Sphere([4.6, 4.4, 0.8], 0.3, matte, Sail.Color.BLACK, reverseNormal=True), [3, 3, 3]))
    return _finish(s, W, H)


def single_rows(W, H):
    """n = 1, tn = 2, ln = 1 (the shader's i/(n-1) row coordinates are 0/0): one emissive sphere that is also
    the only area light's geometry; the camera sits inside it."""
    s = Sail.Scene()
    s.add(Sail.AreaLight(Sail.Sphere([2.78, 2.73, 0.0], 9.0, Sail.Matte(0.6, 10), Sail.UniformColor([0.7, 0.8, 0.9]), reverseNormal=True), [0.4, 0.4, 0.4]))
    return _finish(s, W, H)


def single_object_no_light(W, H):
    """n = 1, ln = 0: a lone Cornell box (material row 0 = its own, Q3); every path dies on black walls."""
    s = Sail.Scene()
    s.add(Sail.Cornellbox())
    return _finish(s, W, H)


MATRIX = {
    "oren_nayar": oren_nayar,
    "rough_glass": rough_glass,
    "textures": textures,
    "light_shapes": light_shapes,
    "reverse_normals": reverse_normals,
    "single_rows": single_rows,
    "single_object_no_light": single_object_no_light,
}


def random_scene(seed, W, H):
    """A seeded random scene over the whole plugin surface: 3-9 objects of random shape kind, material (incl. Oren-Nayar, rough and
    smooth glass, isotropic / anisotropic metal), texture, emission and reverseNormal inside a Cornell box, 0-3 lights of random kind
    (area lights over sphere / rectangle / disk / a stub shape).  Structure fuzz for the per-scene kernel specialisation."""
    import numpy as np
    rng = np.random.default_rng(seed)

    def pos(lo=0.6, hi=4.8):
        return [float(rng.uniform(lo, hi)), float(rng.uniform(0.0, 3.0)), float(rng.uniform(0.4, 4.8))]

    def col():
        return [float(v) for v in rng.uniform(0.1, 1.0, 3)]

    def material():
        k = int(rng.integers(0, 7))
        return [Sail.Matte(0.7), Sail.Matte(0.8, float(rng.uniform(5, 60))), Sail.Mirror(float(rng.uniform(0.5, 1.0))),
                Sail.Metal(0, float(rng.uniform(0.02, 0.3)), float(rng.uniform(0.02, 0.3))), Sail.Metal(float(rng.uniform(0.05, 0.2))),
                Sail.Glass(1, 1, float(rng.uniform(1.2, 1.8))),
                Sail.Glass(0.9, 0.9, 1.5, float(rng.uniform(0.05, 0.3)), float(rng.uniform(0.05, 0.3)))][k]

    def texture():
        k = int(rng.integers(0, 8))
        return [Sail.UniformColor(col()), Sail.UniformColor(col()), Sail.Checkerboard(float(rng.uniform(0.05, 0.3)), 0.02),
                Sail.Checkerboard2(col(), col(), float(rng.uniform(0.05, 0.3))), Sail.Bilerp(col(), col(), col(), col()),
                Sail.Mix(col(), col(), float(rng.uniform(0, 1))), Sail.Scale(col(), col()), Sail.UV()][k]

    def shape(mat, tex, emission=(0, 0, 0), rev=False):
        k = int(rng.integers(0, 8))
        p = pos()
        if k == 0:
            return Sail.Cube(p, [p[0] + float(rng.uniform(0.3, 1.2)), p[1] + float(rng.uniform(0.3, 1.2)), p[2] + float(rng.uniform(0.3, 1.2))], mat, tex, emission, rev)
        if k == 1:
            return Sail.Sphere(p, float(rng.uniform(0.3, 1.0)), mat, tex, emission, rev)
        if k == 2:
            return Sail.Rectangle(p, [p[0] + float(rng.uniform(0.5, 1.5)), p[1] + float(rng.uniform(0.0, 1.0)), p[2] + float(rng.uniform(0.2, 1.0))], mat, tex, emission, rev)
        if k == 3:
            return Sail.Cone(p, float(rng.uniform(0.5, 1.5)), float(rng.uniform(0.3, 0.8)), mat, tex, emission, rev)
        if k == 4:
            return Sail.Cylinder(p, float(rng.uniform(0.5, 1.5)), float(rng.uniform(0.3, 0.7)), mat, tex, emission, rev)
        if k == 5:
            return Sail.Disk(p, float(rng.uniform(0.4, 0.9)), float(rng.uniform(0.0, 0.3)), mat, tex, emission, rev)
        if k == 6:
            return Sail.Hyperboloid(p, [0.4, -0.3, -0.6], [float(rng.uniform(0.3, 0.6)), 0.3, float(rng.uniform(0.4, 0.8))], mat, tex, emission, rev)
        return Sail.Paraboloid(p, 0, float(rng.uniform(0.5, 1.2)), float(rng.uniform(0.3, 0.7)), mat, tex, emission, rev)

    s = Sail.Scene()
    s.add(Sail.Cornellbox())
    for _ in range(int(rng.integers(2, 9))):
        emission = [float(v) for v in rng.uniform(0, 4, 3)] if rng.uniform() < 0.2 else (0, 0, 0)
        s.add(shape(material(), texture(), emission, bool(rng.uniform() < 0.15)))
    matte = Sail.Matte(0.7)
    for _ in range(int(rng.integers(0, 4))):
        k = int(rng.integers(0, 5))
        e = [float(v) for v in rng.uniform(0.5, 6, 3)]
        if k == 0:
            s.add(Sail.AreaLight(Sail.Rectangle([2.13, 5.48, 2.27], [3.43, 5.48, 3.32], matte, Sail.Color.BLACK), e))
        elif k == 1:
            s.add(Sail.AreaLight(Sail.Sphere([float(rng.uniform(1, 4.5)), 4.3, float(rng.uniform(1, 4))], 0.4, matte, Sail.Color.BLACK), e))
        elif k == 2:
            s.add(Sail.PointLight([float(rng.uniform(0.5, 5)), 4.6, float(rng.uniform(0.5, 5))], e))
        elif k == 3:
            s.add(Sail.SpotLight([float(rng.uniform(1, 4.5)), 5.2, float(rng.uniform(1, 4))], float(rng.uniform(20, 50)), 10, [10 * v for v in e]))
        else:
            s.add(Sail.AreaLight(shape(matte, Sail.Color.BLACK), e))  # any shape kind: a real sampler or a stub (Q7)
    return _finish(s, W, H)
