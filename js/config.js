// Lowers scene.tracerConfig() / scene.rendererConfig() (src/scene/scene.js:70-118) to the sail_config POD of
// include/sail_b200.h.  Category ids are those of src/shader/const/define.glsl:18-44.
import box from '../shader/filter/box';
import gaussian from '../shader/filter/gaussian';
import mitchell from '../shader/filter/mitchell';
import sinc from '../shader/filter/sinc';
import triangle from '../shader/filter/triangle';

const SHAPES = {cube: 1, sphere: 2, rectangle: 3, cone: 4, cylinder: 5, disk: 6, hyperboloid: 7, paraboloid: 8, cornellbox: 9};
const MATERIALS = {matte: 1, mirror: 2, metal: 3, glass: 4};
const TEXTURES = {checkerboard: 5, checkerboard2: 7, bilerp: 8, mixf: 9, scale: 10, uvf: 11};
const LIGHTS = {area: 0, point: 1, spot: 2};
const FILTERS = {color: 0, gamma: 1, box: 2, gaussian: 3, mitchell: 4, sinc: 5, triangle: 6, normal: 7, position: 8,
                 wavelet: 9, tonemapping: 10};
const WINDOWS = {box, gaussian, mitchell, sinc, triangle};

function mask(list, ids) {
    let m = 0;
    for (let p of list) m |= 1 << ids[p.name];
    return m >>> 0;
}

// The reference bakes `float windowWeightTable[16] = float[16](...)` into the shader source
// (src/shader/filter/box.js:13-17); the same generator output is parsed back into numbers here.
function windowWeights(name, params) {
    let text = WINDOWS[name].param(params);
    let m = /float\[16\]\(([^)]*)\)/.exec(text);
    return m[1].split(',').map(parseFloat);
}

export function lowerConfig(tracerConfig, rendererConfig, maxBounces = 5) {
    let f = rendererConfig.filter, r = f.getParam('r') || [0, 0];
    let cfg = {
        shapeMask: mask(tracerConfig.shape, SHAPES), materialMask: mask(tracerConfig.material, MATERIALS),
        textureMask: mask(tracerConfig.texture, TEXTURES), lightMask: mask(tracerConfig.light, LIGHTS),
        traceId: 0, maxBounces: maxBounces, accumMode: 0, kernel: 2,
        filterId: FILTERS[f.name], filterGammaC: (f.getParam('c') || [2.2])[0], filterRadius: [r[0], r[1]],
        filterWeights: new Array(16).fill(0)
    };
    if (f.name in WINDOWS) cfg.filterWeights = windowWeights(f.name, f);
    return cfg;
}
