"""Lights of the Sail host API (mirror of /root/reference/src/scene/light.js:7-107).

``gen()`` returns one 18-float lights row (``LIGHTS_LENGTH``, webgl.js:138): category id
(define.glsl:28-30), parameters, then the emission appended by ``Light.gen`` (light.js:20-27).
"""
import math

LIGHTS_LENGTH = 18


class Light:
    pluginName = ''

    def __init__(self, emission):
        self.emission = [float(c) for c in emission]

    def _gen(self, data):
        data = [float(v) for v in data] + list(self.emission)
        return data + [0.0] * (LIGHTS_LENGTH - len(data))


class GeometryLight(Light):
    """light.js:30-56: a light carried by a shape that is also pushed into ``scene.objects``."""

    def __init__(self, geometry, emission):
        super().__init__(emission)
        geometry.emission = [float(c) for c in emission]
        self._geometry = geometry
        self.index = None

    @property
    def geometry(self):
        # light.js:41-43: only visible once the scene has assigned an index
        return self._geometry if self.index is not None else None

    def getGeometry(self, index):
        self.index = index
        return self._geometry

    def _gen(self, data):
        if self.index is None:
            raise RuntimeError("can't find index of AreaLight's geometry")  # light.js:52
        return super()._gen(list(data) + [self.index])


class AreaLight(GeometryLight):
    """light.js:58-69.  Row [0, objectIndex, emission(3)]."""

    pluginName = 'area'

    def gen(self):
        return self._gen([0])


class PointLight(Light):
    """light.js:71-87.  Row [1, from(3), emission(3)]."""

    pluginName = 'point'

    def __init__(self, from_, emission):
        super().__init__(emission)
        self.from_ = [float(c) for c in from_]

    def gen(self):
        return self._gen([1, *self.from_])


class SpotLight(Light):
    """light.js:89-107.  Row [2, cosTotalWidth, cosFalloffStart, from(3), emission(3)]."""

    pluginName = 'spot'

    def __init__(self, from_, coneangle, conedelta, emission):
        super().__init__(emission)
        self.cosTotalWidth = math.cos(coneangle / 180 * math.pi)
        self.cosFalloffStart = math.cos((coneangle - conedelta) / 180 * math.pi)
        self.from_ = [float(c) for c in from_]

    def gen(self):
        return self._gen([2, self.cosTotalWidth, self.cosFalloffStart, *self.from_])
