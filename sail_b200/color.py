"""Color helper (mirror of /root/reference/src/core/color.js:7-30)."""
from .texture import UniformColor


class _ColorMeta(type):
    # color.js:15-29 are static getters returning a fresh UniformColor each time.
    @property
    def BLACK(cls):
        return UniformColor([0, 0, 0])

    @property
    def WHITE(cls):
        return UniformColor([1, 1, 1])

    @property
    def GREEN(cls):
        return UniformColor([0, 1, 0])

    @property
    def BLUE(cls):
        return UniformColor([0, 0, 1])

    @property
    def RED(cls):
        return UniformColor([1, 0, 0])


class Color(metaclass=_ColorMeta):
    @staticmethod
    def createTexture(color):
        """color.js:8-13."""
        return UniformColor(list(color))

    #: README.md:49 uses ``Sail.Color.create``; kept as an alias.
    create = createTexture
