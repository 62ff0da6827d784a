"""Plugin selection API (mirror of the public shape of /root/reference/src/shader/generator.js).

The reference concatenates GLSL per plugin name at run time (generator.js:107-123).  Here the
plugin *names* and parameters are kept as the host-visible knobs (``scene.filter = 'gaussian'``,
``scene.filter.addParam('r', 'vec2(2.0,2.0)')``) and are lowered to category bit masks and a POD
parameter block for ``sail_configure`` -- the CUDA kernels are compiled ahead of time.
"""
import re

#: shader.shape.js:16-26, category ids from const/define.glsl:18-26
SHAPE_PLUGINS = {"cube": 1, "sphere": 2, "rectangle": 3, "cone": 4, "cylinder": 5, "disk": 6,
                 "hyperboloid": 7, "paraboloid": 8, "cornellbox": 9}
#: shader.material.js:14-19, define.glsl:32-35
MATERIAL_PLUGINS = {"matte": 1, "mirror": 2, "metal": 3, "glass": 4}
#: shader.texture.js:13-20, define.glsl:38-44
TEXTURE_PLUGINS = {"checkerboard": 5, "checkerboard2": 7, "bilerp": 8, "mixf": 9, "scale": 10, "uvf": 11}
#: shader.light.js:6-10, define.glsl:28-30
LIGHT_PLUGINS = {"area": 0, "point": 1, "spot": 2}
#: shader.trace.js:7-9
TRACE_PLUGINS = {"path": 0}
#: shader.filter.js:18-30 (ids are this library's: include/sail_b200.h SAIL_FILTER_*)
FILTER_PLUGINS = {"color": 0, "gamma": 1, "box": 2, "gaussian": 3, "mitchell": 4, "sinc": 5, "triangle": 6,
                  "normal": 7, "position": 8, "wavelet": 9, "tonemapping": 10}

_FLOAT_RE = re.compile(r"-?\d+\.\d+?")


class PluginParams:
    """generator.js:4-25."""

    def __init__(self, name):
        self.name = name
        self.params = {}

    def addParam(self, name, value):
        self.params[name] = value

    def getParam(self, name):
        """generator.js:14-20: the regex ``/-?\\d+\\.\\d+?/g`` keeps ONE fractional digit
        ("2.25" -> 2.2), hence "the format of params' value must be float" (ui/ui.js:42, Q10)."""
        value = self.params.get(name)
        if value is None:
            return None
        return [float(m) for m in _FLOAT_RE.findall(value)]

    def getParamName(self, name, generatorName):
        return ("%s_%s_%s" % (generatorName, self.name, name)).upper()


def query(registry, plugin):
    """generator.js:125-127."""
    return plugin in registry
