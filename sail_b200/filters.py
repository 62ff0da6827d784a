"""Host side of the display filters: the 16-entry window weight tables that the reference's JS
bakes into the render shader (mirror of /root/reference/src/shader/filter/{box,gaussian,mitchell,
sinc,triangle}.js and shader.filter.js:31 ``windowWidth = 4``)."""
import math

from .generator import FILTER_PLUGINS

WINDOW_WIDTH = 4


def _box(px, py, r, params):
    return 1.0  # box.js:4-6


def _gaussian(px, py, r, params):
    # gaussian.js:4-6,27-28
    alpha = params["alpha"][0]
    g = lambda d, expv: max(0.0, math.exp(-alpha * d * d) - expv)
    return g(px, math.exp(-alpha * r[0] * r[0])) * g(py, math.exp(-alpha * r[1] * r[1]))


def _mitchell1(x, B, C):
    # mitchell.js:4-14
    x = abs(2 * x)
    if x > 1:
        return ((-B - 6 * C) * x * x * x + (6 * B + 30 * C) * x * x + (-12 * B - 48 * C) * x + (8 * B + 24 * C)) * (1.0 / 6.0)
    return ((12 - 9 * B - 6 * C) * x * x * x + (-18 + 12 * B + 6 * C) * x * x + (6 - 2 * B)) * (1.0 / 6.0)


def _mitchell(px, py, r, params):
    # mitchell.js:36-37; `b`, `c` are 1-element arrays coerced to numbers by JS arithmetic
    b, c = params["b"][0], params["c"][0]
    return _mitchell1(px / r[0], b, c) * _mitchell1(py / r[1], b, c)


def _sinc1(x):
    x = abs(x)
    if x < 1e-5:
        return 1.0
    return math.sin(math.pi * x) / (math.pi * x)


def _windowed_sinc(x, radius, tau):
    # sinc.js:10-15
    x = abs(x)
    if x > radius:
        return 0.0
    return _sinc1(x) * _sinc1(x / tau)


def _sinc(px, py, r, params):
    tau = params["tau"][0]  # sinc.js:21: 1-element array, coerced
    return _windowed_sinc(px, r[0], tau) * _windowed_sinc(py, r[1], tau)


def _triangle(px, py, r, params):
    # triangle.js:4-6,26-27
    return max(0.0, r[0] - px) * max(0.0, r[1] - py)


_WINDOW = {"box": _box, "gaussian": _gaussian, "mitchell": _mitchell, "sinc": _sinc, "triangle": _triangle}
_WINDOW_NEEDS = {"box": (), "gaussian": ("alpha",), "mitchell": ("b", "c"), "sinc": ("tau",), "triangle": ()}


def window_weight_table(name, plugin_params):
    """Returns (radius[2], weights[16]) for a window filter, indexed [i*4+j] like the JS loops
    (box.js:19-32); the shader then reads it with ``i*j+j`` (window.glsl:38)."""
    r = plugin_params.getParam("r")
    if not r or len(r) < 2:
        raise ValueError("filter '%s' needs param r = 'vec2(x.x,y.y)'" % name)
    params = {}
    for key in _WINDOW_NEEDS[name]:
        v = plugin_params.getParam(key)
        if not v:
            raise ValueError("filter '%s' needs param %s" % (name, key))
        params[key] = v
    weights = []
    for i in range(WINDOW_WIDTH):
        for j in range(WINDOW_WIDTH):
            px = (j + 0.5) * r[0] / WINDOW_WIDTH
            py = (i + 0.5) * r[1] / WINDOW_WIDTH
            weights.append(float(_WINDOW[name](px, py, r, params)))
    return [r[0], r[1]], weights


def lower_filter(plugin_params):
    """PluginParams -> dict(filter_id, gamma_c, radius, weights) for sail_configure."""
    name = plugin_params.name
    out = {"filter_id": FILTER_PLUGINS[name], "gamma_c": 2.2, "radius": [0.0, 0.0], "weights": [0.0] * 16}
    if name in _WINDOW:
        out["radius"], out["weights"] = window_weight_table(name, plugin_params)
    elif name == "gamma":
        c = plugin_params.getParam("c")
        if not c:
            raise ValueError("filter 'gamma' needs param c")  # FILTER_GAMMA_C would be undefined -> link failure
        out["gamma_c"] = c[0]
    elif name == "wavelet":
        r = plugin_params.getParam("r")
        if not r or len(r) < 2:
            raise ValueError("filter 'wavelet' needs param r = 'vec2(x.x,y.y)'")
        out["radius"] = [r[0], r[1]]
    return out
