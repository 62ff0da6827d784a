"""Materials of the Sail host API (mirror of /root/reference/src/scene/material.js:14-124).

Each ``gen()`` returns one 16-float texParams row (``ShaderProgram.TEXPARAMS_LENGTH``,
/root/reference/src/core/webgl.js:137-139), zero padded like ``Material.gen`` (material.js:25-29).
"""
import math

TEXPARAMS_LENGTH = 16

#: material.js:87-88 -- default conductor eta / k (copper-like).
_DEFAULT_ETA = (9.530817595377695, 6.635831967341377, 4.47513354108444)
_DEFAULT_K = (13.028170336874789, 8.112634272577575, 5.502811570992323)


def _pad(data, length=TEXPARAMS_LENGTH):
    data = [float(v) for v in data]
    if len(data) > length:
        raise ValueError("row longer than %d floats" % length)
    return data + [0.0] * (length - len(data))


class Material:
    """material.js:14-30."""

    pluginName = ''

    def gen(self):  # pragma: no cover - abstract
        raise NotImplementedError


class Matte(Material):
    """material.js:32-59.  Row: [1, kd, sigma(deg), A, B]; Oren-Nayar A/B from sigma (:44-49)."""

    pluginName = 'matte'

    def __init__(self, kd=1, sigma=0):
        if kd <= 0:
            kd = 1
        self.kd = kd
        self.sigma = sigma
        self.A = 0.0
        self.B = 0.0
        if self.sigma != 0:
            s = sigma * math.pi / 180
            s2 = s * s
            self.A = 1.0 - (s2 / (2.0 * (s2 + 0.33)))
            self.B = 0.45 * s2 / (s2 + 0.09)

    def gen(self):
        return _pad([1, self.kd, self.sigma, self.A, self.B])


class Mirror(Material):
    """material.js:61-78.  Row: [2, kr]."""

    pluginName = 'mirror'

    def __init__(self, kr=1.0):
        if kr <= 0:
            kr = 0.5
        self.kr = kr

    def gen(self):
        return _pad([2, self.kr])


class Metal(Material):
    """material.js:80-102.  Row: [3, uroughness, vroughness, eta(3), k(3)].

    ``roughnessToAlpha`` (material.js:7-12) is never applied by the reference: the raw roughness is
    used as the Trowbridge-Reitz alpha (SURVEY.md 8(a) "metal").
    """

    pluginName = 'metal'

    def __init__(self, roughness=0.01, uroughness=0, vroughness=0, eta=None, k=None):
        self.uroughness = roughness if uroughness == 0 else uroughness
        self.vroughness = roughness if vroughness == 0 else vroughness
        self.eta = tuple(eta) if eta is not None else _DEFAULT_ETA
        self.k = tuple(k) if k is not None else _DEFAULT_K

    def gen(self):
        return _pad([3, self.uroughness, self.vroughness, *self.eta, *self.k])


class Glass(Material):
    """material.js:104-124.  Row: [4, kr, kt, eta, uroughness, vroughness]."""

    pluginName = 'glass'

    def __init__(self, kr=1, kt=1, eta=None, uroughness=0, vroughness=0):
        if eta is None:
            eta = float('nan')  # material.js:105 has no default: `undefined` becomes NaN in the Float32Array
        self.kr = kr
        self.kt = kt
        self.eta = eta
        self.uroughness = uroughness
        self.vroughness = vroughness

    def gen(self):
        return _pad([4, self.kr, self.kt, self.eta, self.uroughness, self.vroughness])
