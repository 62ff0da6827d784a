"""Shape classes of the Sail host API (mirror of /root/reference/src/scene/geometry.js:27-632).

``gen(texparamID)`` returns the 18-float objects row of SURVEY.md appendix A: category id, shape
parameters, then ``reverseNormal, matRow, texRow, emission.rgb`` appended by ``Object3D.gen``
(geometry.js:82-90), zero padded to ``OBJECTS_LENGTH`` = 18 (webgl.js:137).
``genTexparams()`` returns material row ++ texture row (geometry.js:75-80).
``boundbox()`` is the picking bound of the interaction layer (pickup.js:41-62): ``{'min', 'max'}`` with any
pending ``temporaryTranslation`` NOT applied (as in the reference), or ``False`` for shapes that cannot be
picked (geometry.js:51-53: Object3D, hence Cornellbox).  The per-shape CPU ``intersect(ray)`` of the reference is
replaced by the device query behind ``sail_b200.pickup.Pickup`` (sail_intersect_rays).
"""
import math

from .material import Matte
from .texture import UniformColor

OBJECTS_LENGTH = 18


def _vec(v):
    return [float(c) for c in v]


class Object3D:
    """geometry.js:27-91."""

    pluginName = ''

    def __init__(self, material, texture, emission=(0, 0, 0), reverseNormal=False):
        self.material = material
        self.texture = texture
        self.emission = _vec(emission if emission is not None else (0, 0, 0))
        self.reverseNormal = 1 if reverseNormal else 0
        self.texparamID = 0
        self.temporaryTranslation = [0.0, 0.0, 0.0]
        self.light = any(c != 0 for c in self.emission)

    def boundbox(self):
        """geometry.js:51-53."""
        return False

    def temporaryTranslate(self, vector):
        self.temporaryTranslation = _vec(vector)

    def translate(self):
        self.temporaryTranslation = [0.0, 0.0, 0.0]

    def _moved(self, p):
        return [p[i] + self.temporaryTranslation[i] for i in range(3)]

    def genTexparams(self):
        return list(self.material.gen()) + list(self.texture.gen())

    def _gen(self, data, texparamID):
        data = [float(v) for v in data]
        data += [float(self.reverseNormal), float(texparamID), float(texparamID + 1), *self.emission]
        if len(data) > OBJECTS_LENGTH:
            raise ValueError("objects row longer than %d floats" % OBJECTS_LENGTH)
        return data + [0.0] * (OBJECTS_LENGTH - len(data))

    def _begin_gen(self, texparamID):
        if texparamID is None:
            texparamID = self.texparamID
        self.texparamID = texparamID
        return texparamID


class Cube(Object3D):
    """geometry.js:93-139.  Row: [1, min(3), max(3), rev, mat, tex, emission(3)]."""

    pluginName = 'cube'

    def __init__(self, min, max, material, texture, emission=(0, 0, 0), reverseNormal=False):
        super().__init__(material, texture, emission, reverseNormal)
        self.min = _vec(min)
        self.max = _vec(max)

    def translate(self):
        self.min = self._moved(self.min)
        self.max = self._moved(self.max)
        super().translate()

    def boundbox(self):
        """geometry.js:103-108."""
        return {'min': list(self.min), 'max': list(self.max)}

    def gen(self, texparamID=None):
        texparamID = self._begin_gen(texparamID)
        self.translate()
        return self._gen([1, *self.min, *self.max], texparamID)


class Sphere(Object3D):
    """geometry.js:141-190.  Row: [2, c(3), r, rev, mat, tex, emission(3)]."""

    pluginName = 'sphere'

    def __init__(self, c, r, material, texture, emission=(0, 0, 0), reverseNormal=False):
        super().__init__(material, texture, emission, reverseNormal)
        self.c = _vec(c)
        self.r = float(r)

    def translate(self):
        self.c = self._moved(self.c)
        super().translate()

    def boundbox(self):
        """geometry.js:152-158."""
        return {'min': [c - self.r for c in self.c], 'max': [c + self.r for c in self.c]}

    def gen(self, texparamID=None):
        texparamID = self._begin_gen(texparamID)
        self.translate()
        return self._gen([2, *self.c, self.r], texparamID)


class Rectangle(Object3D):
    """geometry.js:192-266.  Row: [3, min(3), max(3), rev, mat, tex, emission(3)]."""

    pluginName = 'rectangle'

    def __init__(self, min, max, material, texture, emission=(0, 0, 0), reverseNormal=False):
        super().__init__(material, texture, emission, reverseNormal)
        self.min = _vec(min)
        self.max = _vec(max)

    def translate(self):
        self.min = self._moved(self.min)
        self.max = self._moved(self.max)
        super().translate()

    def boundbox(self):
        """geometry.js:202-222: a flat axis is padded by 0.05 on both sides."""
        mn, mx = list(self.min), list(self.max)
        for a in range(3):
            if self.max[a] == self.min[a]:
                mx[a] += 0.05
                mn[a] -= 0.05
        return {'min': mn, 'max': mx}

    def gen(self, texparamID=None):
        texparamID = self._begin_gen(texparamID)
        self.translate()
        return self._gen([3, *self.min, *self.max], texparamID)


class _Positioned(Object3D):
    def translate(self):
        self.position = self._moved(self.position)
        super().translate()


class Cone(_Positioned):
    """geometry.js:268-332.  Row: [4, position(3), height, radius, rev, mat, tex, emission(3)]."""

    pluginName = 'cone'

    def __init__(self, position, height, radius, material, texture, emission=(0, 0, 0), reverseNormal=False):
        super().__init__(material, texture, emission, reverseNormal)
        self.position = _vec(position)
        self.height = float(height)
        self.radius = float(radius)

    def boundbox(self):
        """geometry.js:279-284."""
        p = self.position
        return {'min': [p[0] - self.radius, p[1], p[2] - self.radius],
                'max': [p[0] + self.radius, p[1] + self.height, p[2] + self.radius]}

    def gen(self, texparamID=None):
        texparamID = self._begin_gen(texparamID)
        self.translate()
        return self._gen([4, *self.position, self.height, self.radius], texparamID)


class Cylinder(_Positioned):
    """geometry.js:334-397.  Row: [5, position(3), height, radius, rev, mat, tex, emission(3)]."""

    pluginName = 'cylinder'

    def __init__(self, position, height, radius, material, texture, emission=(0, 0, 0), reverseNormal=False):
        super().__init__(material, texture, emission, reverseNormal)
        self.position = _vec(position)
        self.height = float(height)
        self.radius = float(radius)

    def boundbox(self):
        """geometry.js:345-350."""
        p = self.position
        return {'min': [p[0] - self.radius, p[1], p[2] - self.radius],
                'max': [p[0] + self.radius, p[1] + self.height, p[2] + self.radius]}

    def gen(self, texparamID=None):
        texparamID = self._begin_gen(texparamID)
        self.translate()
        return self._gen([5, *self.position, self.height, self.radius], texparamID)


class Disk(_Positioned):
    """geometry.js:399-449.  Row: [6, position(3), radius, innerRadius, rev, mat, tex, emission(3)]."""

    pluginName = 'disk'

    def __init__(self, position, radius, innerRadius, material, texture, emission=(0, 0, 0), reverseNormal=False):
        super().__init__(material, texture, emission, reverseNormal)
        self.position = _vec(position)
        self.radius = float(radius)
        self.innerRadius = float(innerRadius)

    def boundbox(self):
        """geometry.js:410-415."""
        p = self.position
        return {'min': [p[0] - self.radius, p[1] - 0.05, p[2] - self.radius],
                'max': [p[0] + self.radius, p[1] + 0.05, p[2] + self.radius]}

    def gen(self, texparamID=None):
        texparamID = self._begin_gen(texparamID)
        self.translate()
        return self._gen([6, *self.position, self.radius, self.innerRadius], texparamID)


class Hyperboloid(_Positioned):
    """geometry.js:451-540.  Row: [7, position(3), p1(3), p2(3), ah, ch, rev, mat, tex, emission(3)].

    ``ah``/``ch`` are the implicit-form coefficients fitted on the host (geometry.js:467-486).
    """

    pluginName = 'hyperboloid'

    def __init__(self, position, p1, p2, material, texture, emission=(0, 0, 0), reverseNormal=False):
        super().__init__(material, texture, emission, reverseNormal)
        self.position = _vec(position)
        self.p1 = _vec(p1)
        self.p2 = _vec(p2)
        r1 = math.sqrt(p1[0] * p1[0] + p1[1] * p1[1])
        r2 = math.sqrt(p2[0] * p2[0] + p2[1] * p2[1])
        self.rMax = max(r1, r2)
        self.zMin = min(p1[2], p2[2])
        self.zMax = max(p1[2], p2[2])
        if self.p2[2] == 0:
            self.p1, self.p2 = self.p2, self.p1
        pp = list(self.p1)
        n = 0
        while True:
            pp = [pp[i] + (self.p2[i] - self.p1[i]) * 2 for i in range(3)]
            xy1 = pp[0] * pp[0] + pp[1] * pp[1]
            xy2 = self.p2[0] * self.p2[0] + self.p2[1] * self.p2[1]
            z2 = self.p2[2] * self.p2[2]
            try:
                self.ah = (1 / xy1 - (pp[2] * pp[2]) / (xy1 * z2)) / (1 - (xy2 * pp[2] * pp[2]) / (xy1 * z2))
                self.ch = (self.ah * xy2 - 1) / z2
            except ZeroDivisionError:
                self.ah = float('inf')
                self.ch = float('nan')
            n += 1
            # geometry.js:481: `while (!isFinite(ah) || isNaN(ah) && n<20)`; bounded here so that a
            # degenerate input raises (geometry.js:483-485) instead of spinning forever.
            if not (not math.isfinite(self.ah) and n < 20):
                break
        if not math.isfinite(self.ah):
            raise ValueError("the p1,p2 of hyperboloid is illegal")

    def boundbox(self):
        """geometry.js:489-494."""
        p = self.position
        return {'min': [p[0] - self.rMax, p[1] + self.zMin, p[2] - self.rMax],
                'max': [p[0] + self.rMax, p[1] + self.zMax, p[2] + self.rMax]}

    def gen(self, texparamID=None):
        texparamID = self._begin_gen(texparamID)
        self.translate()
        return self._gen([7, *self.position, *self.p1, *self.p2, self.ah, self.ch], texparamID)


class Paraboloid(_Positioned):
    """geometry.js:542-608.  Row: [8, position(3), z0, z1, radius, rev, mat, tex, emission(3)]."""

    pluginName = 'paraboloid'

    def __init__(self, position, z0, z1, radius, material, texture, emission=(0, 0, 0), reverseNormal=False):
        super().__init__(material, texture, emission, reverseNormal)
        self.position = _vec(position)
        self.z0 = float(z0)
        self.z1 = float(z1)
        self.radius = float(radius)

    def boundbox(self):
        """geometry.js:557-562 (its zMin/zMax are z0/z1 sorted, geometry.js:549-550)."""
        p = self.position
        zMin, zMax = min(self.z0, self.z1), max(self.z0, self.z1)
        return {'min': [p[0] - self.radius, p[1] + zMin, p[2] - self.radius],
                'max': [p[0] + self.radius, p[1] + zMax, p[2] + self.radius]}

    def gen(self, texparamID=None):
        texparamID = self._begin_gen(texparamID)
        self.translate()
        return self._gen([8, *self.position, self.z0, self.z1, self.radius], texparamID)


class Cornellbox(Object3D):
    """geometry.js:610-632.  Row: [9, min(3), max(3), rev, mat, tex, emission(3)].

    The shader reads its material row from column 7 (= reverseNormal = 0), i.e. texParams row 0
    (SURVEY.md Q3); the ``Matte(1)`` given here is serialised but never used.
    """

    pluginName = 'cornellbox'

    def __init__(self, min=(0, 0, -5), max=(5.560, 5.488, 5.592)):
        super().__init__(Matte(1), UniformColor([0, 0, 0]))
        self.min = _vec(min)
        self.max = _vec(max)

    def scale(self, k):
        self.min = [c * k for c in self.min]
        self.max = [c * k for c in self.max]

    def gen(self, texparamID=None):
        texparamID = self._begin_gen(texparamID)
        return self._gen([9, *self.min, *self.max], texparamID)
