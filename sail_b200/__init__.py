"""sail_b200 -- host-side mirror of the ``Sail`` namespace (/root/reference/index.js:15-45) over
libsail_b200.so, the sm_100a path-tracing core (include/sail_b200.h).  Only what the hot path needs."""
from .camera import Camera
from .color import Color
from .generator import PluginParams
from .geometry import Cone, Cornellbox, Cube, Cylinder, Disk, Hyperboloid, Paraboloid, Rectangle, Sphere
from .light import AreaLight, PointLight, SpotLight
from .material import Glass, Matte, Metal, Mirror
from .scene import Scene
from .schedule import FrameSchedule
from .texture import UV, Bilerp, Checkerboard, Checkerboard2, Mix, Scale, UniformColor


def __getattr__(name):
    # Renderer / Tracer need the native library; import lazily so the scene API works without it.
    if name in ("Renderer", "Tracer"):
        from . import renderer
        return getattr(renderer, name)
    if name in ("Pickup", "Ray"):  # index.js:19 exports the interaction layer too
        from . import pickup
        return getattr(pickup, name)
    raise AttributeError(name)
