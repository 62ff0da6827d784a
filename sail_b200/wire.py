"""Scene serialisation into the reference wire format (mirror of ``Tracer.update`` /
``Tracer.updateObjects``, /root/reference/src/core/tracer.js:25-60): three flat float32 tables,
objects 18 x n, texParams 16 x tn (material row, texture row per object), lights 18 x ln."""
import numpy as np

from .geometry import OBJECTS_LENGTH
from .light import LIGHTS_LENGTH
from .material import TEXPARAMS_LENGTH


def serialize_scene(scene):
    objects, texparams, lights = [], [], []
    for ob in scene.objects:
        objects.extend(ob.gen(len(texparams) // TEXPARAMS_LENGTH))  # tracer.js:47
        texparams.extend(ob.genTexparams())                          # tracer.js:48
    for light in scene.lights:
        lights.extend(light.gen())                                   # tracer.js:50-52
    return (np.asarray(objects, dtype=np.float32).reshape(-1, OBJECTS_LENGTH),
            np.asarray(texparams, dtype=np.float32).reshape(-1, TEXPARAMS_LENGTH),
            np.asarray(lights, dtype=np.float32).reshape(-1, LIGHTS_LENGTH))


def serialize_objects(scene):
    """tracer.js:25-31: objects only, re-using each object's stored texparamID."""
    objects = []
    for ob in scene.objects:
        objects.extend(ob.gen())
    return np.asarray(objects, dtype=np.float32).reshape(-1, OBJECTS_LENGTH)
