"""Procedural textures of the Sail host API (mirror of /root/reference/src/scene/texture.js:7-170).

Each ``gen()`` returns one 16-float texParams row; column 0 is the texture category id of
/root/reference/src/shader/const/define.glsl:38-44.
"""
from .material import _pad


class Texture:
    """texture.js:7-23."""

    pluginName = ''


class UniformColor(Texture):
    """texture.js:25-41.  Row [0, r, g, b].  ``pluginName`` is undefined in the reference (handled
    inline by getSurfaceColor, shader.texture.js:24)."""

    pluginName = None

    def __init__(self, color):
        self.color = tuple(float(c) for c in color)

    def gen(self):
        return _pad([0, *self.color])


class Checkerboard(Texture):
    """texture.js:43-63.  Row [5, size, lineWidth]."""

    pluginName = 'checkerboard'

    def __init__(self, size=0.1, lineWidth=0.01):
        if size <= 0:
            size = 0.3
        if lineWidth < 0:
            lineWidth = 0.03
        self.size = size
        self.lineWidth = lineWidth

    def gen(self):
        return _pad([5, self.size, self.lineWidth])


class Checkerboard2(Texture):
    """texture.js:65-86.  Row [7, color1(3), color2(3), size]."""

    pluginName = 'checkerboard2'

    def __init__(self, color1=(1, 1, 1), color2=(0, 0, 0), size=0.1):
        self.color1 = tuple(color1)
        self.color2 = tuple(color2)
        self.size = size

    def gen(self):
        return _pad([7, *self.color1, *self.color2, self.size])


class Bilerp(Texture):
    """texture.js:88-111.  Row [8, c00, c01, c10, c11].  The reference shader for it does not compile
    (SURVEY.md Q12); the kernels implement the evident bilinear blend."""

    pluginName = 'bilerp'

    def __init__(self, color00, color01, color10, color11):
        self.color00 = tuple(color00)
        self.color01 = tuple(color01)
        self.color10 = tuple(color10)
        self.color11 = tuple(color11)

    def gen(self):
        return _pad([8, *self.color00, *self.color01, *self.color10, *self.color11])


class Mix(Texture):
    """texture.js:113-134.  Row [9, color1, color2, amount]."""

    pluginName = 'mixf'

    def __init__(self, color1, color2, amount):
        self.color1 = tuple(color1)
        self.color2 = tuple(color2)
        self.amount = amount

    def gen(self):
        return _pad([9, *self.color1, *self.color2, self.amount])


class Scale(Texture):
    """texture.js:136-155.  Row [10, color1, color2]."""

    pluginName = 'scale'

    def __init__(self, color1, color2):
        self.color1 = tuple(color1)
        self.color2 = tuple(color2)

    def gen(self):
        return _pad([10, *self.color1, *self.color2])


class UV(Texture):
    """texture.js:157-170.  Row [11]."""

    pluginName = 'uvf'

    def gen(self):
        return _pad([11])
