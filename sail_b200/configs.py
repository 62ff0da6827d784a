"""The five workloads of BASELINE.json ``configs`` as Sail scene scripts.

Each builder returns ``(scene, width, height, frames, max_bounces)``.  Scene constants follow
/root/reference/README.md:36-55 (config 1) and /root/reference/ui/ui.js:10-44 (config 2); configs 3-5
are the extensions fixed in SURVEY.md 8(d).
"""
from .camera import Camera
from .color import Color
from .geometry import Cone, Cornellbox, Cube, Cylinder, Disk, Hyperboloid, Paraboloid, Rectangle, Sphere
from .light import AreaLight, PointLight, SpotLight
from .material import Glass, Matte, Metal, Mirror
from .scene import Scene
from .texture import UV, Checkerboard, Checkerboard2, Mix, UniformColor


def _camera(width, height):
    cam = Camera([2.78, 2.73, -6], [2.78, 2.73, 2.79])  # README.md:40, ui/ui.js:12-14
    cam.makePerspective(55, width / height, 1, 100)     # aspect generalised from 1 (camera.js:16, Q9)
    return cam


def config1(width=512, height=512):
    """README Cornell box: area-light cube, Cornell box walls, mirror sphere; no NEE lights."""
    scene = Scene()
    scene.add(Cube([2.13, 5.487, 2.27], [3.43, 5.488, 3.32], Matte(0.7), UniformColor([0, 0, 0]), [8, 8, 8]))
    scene.add(Cornellbox([0, 0, -7], [5.560, 5.488, 5.592]))
    scene.add(Sphere([2, 1.25, 2.70], 1.2, Mirror(1.0), UniformColor([1, 1, 1])))
    scene.add(_camera(width, height))
    scene.filter = 'color'
    return scene, width, height, 64, 5


def config2(width=1920, height=1080):
    """Materials demo: mirror / glass / metal / matte spheres under an area light."""
    scene = Scene()
    matte = Matte(0.7)
    scene.add(AreaLight(Rectangle([2.13, 5.48, 2.27], [3.43, 5.48, 3.32], matte, Color.BLACK), [1, 1, 1]))
    scene.add(Cornellbox())
    scene.add(Sphere([1.5, 1.25, 2.70], 1.2, Mirror(1.0), Color.WHITE))
    scene.add(Sphere([3.9, 1.25, 1.70], 1.2, Glass(1, 1, 1.5), Color.WHITE))
    scene.add(Sphere([2.78, 0.6, 0.2], 0.6, Metal(0, 0.01, 0.1), Color.WHITE))
    scene.add(Sphere([2.78, 0.7, 4.4], 0.7, Matte(0.7), Color.WHITE))
    scene.add(_camera(width, height))
    scene.filter = 'tonemapping'
    return scene, width, height, 64, 5


def config3(width=1920, height=1080):
    """Quadric/cone/rectangle/disk shapes, checkerboard + linear-blend textures, point+spot+area lights."""
    scene = Scene()
    matte = Matte(0.7)
    scene.add(Cornellbox())
    scene.add(Cone([1.0, 0, 3.6], 1.6, 0.7, matte, Checkerboard(0.1, 0.01)))
    scene.add(Cylinder([4.4, 0, 3.8], 1.2, 0.5, matte, Checkerboard2([1, 1, 1], [0, 0, 0], 0.1)))
    scene.add(Hyperboloid([2.78, 1.0, 3.0], [0.4, -0.3, -0.6], [0.4, 0.3, 0.6], Metal(0, 0.05, 0.05), Color.WHITE))
    scene.add(Paraboloid([1.4, 0, 1.2], 0, 1.0, 0.6, Glass(1, 1, 1.5), Color.WHITE))
    scene.add(Disk([4.2, 0.01, 1.4], 0.8, 0.3, matte, UV()))
    scene.add(Rectangle([0.6, 0.8, 5.5], [4.9, 3.8, 5.5], matte, Mix([0.9, 0.2, 0.2], [0.2, 0.2, 0.9], 0.35)))
    scene.add(AreaLight(Rectangle([2.13, 5.48, 2.27], [3.43, 5.48, 3.32], matte, Color.BLACK), [4, 4, 4]))
    scene.add(PointLight([0.8, 4.6, 0.8], [0.6, 0.6, 0.6]))
    scene.add(SpotLight([4.6, 5.2, 2.0], 35, 10, [30, 30, 30]))
    scene.add(_camera(width, height))
    scene.filter = 'tonemapping'
    return scene, width, height, 256, 5


def config4(width=3840, height=2160):
    """Config-1 Cornell box at 4K, 1024 spp, sample-sharded across GPUs."""
    scene, _, _, _, bounces = config1(width, height)
    return scene, width, height, 1024, bounces


def config5(width=1920, height=1080, filter_name='box'):
    """Filter sweep on the glass + metal scene, 16 bounces, tile-sharded."""
    scene = Scene()
    matte = Matte(0.7)
    scene.add(AreaLight(Rectangle([2.13, 5.48, 2.27], [3.43, 5.48, 3.32], matte, Color.BLACK), [1, 1, 1]))
    scene.add(Cornellbox())
    scene.add(Sphere([3.9, 1.25, 1.70], 1.2, Glass(1, 1, 1.5), Color.WHITE))
    scene.add(Sphere([2.78, 0.6, 0.2], 0.6, Metal(0, 0.01, 0.1), Color.WHITE))
    scene.add(_camera(width, height))
    scene.filter = filter_name
    scene.filter.addParam('r', 'vec2(2.0,2.0)')
    scene.filter.addParam('tau', '3.0')
    scene.filter.addParam('alpha', '2.0')
    return scene, width, height, 64, 16


CONFIGS = {1: config1, 2: config2, 3: config3, 4: config4, 5: config5}


def cornell_1080p():
    """The headline workload (BASELINE.json ``metric``): the README Cornell box at 1920x1080."""
    scene, _, _, _, bounces = config1(1920, 1080)
    return scene, 1920, 1080, 64, bounces
