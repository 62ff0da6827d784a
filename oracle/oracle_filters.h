// oracle/oracle_filters.h -- TEST INFRASTRUCTURE ONLY.  Restates the display pass:
//   src/shader/main/vsrender.glsl:4-7, main/fsrender.glsl:4-6 and src/shader/filter/
//   color.glsl:1-4, gamma.glsl:1-8, tonemapping.glsl:1-9, normal.glsl:1-3, position.glsl:1-3,
//   window.glsl:1-44 (box/gaussian/mitchell/sinc/triangle share it; the 16-entry weight table is built
//   on the host by filter/{box,gaussian,mitchell,sinc,triangle}.js and arrives here as floats),
//   wavelet.glsl:1-54.
// The accumulation textures are sampled LINEAR with the default REPEAT wrap (webgl.js:153-156).  The
// bilinear weights are quantised to 8 fractional bits like GPU texture units, so a fetch at a pixel
// centre returns that texel exactly.  512 is generalised to W (x) / H (y) and 1280 to 2.5*W / 2.5*H
// (SURVEY.md Q9).
#ifndef SAIL_ORACLE_FILTERS_H
#define SAIL_ORACLE_FILTERS_H
#include "shader_util.h"

enum { F_COLOR = 0, F_GAMMA = 1, F_BOX = 2, F_GAUSSIAN = 3, F_MITCHELL = 4, F_SINC = 5, F_TRIANGLE = 6, F_NORMAL = 7,
       F_POSITION = 8, F_WAVELET = 9, F_TONEMAPPING = 10 };

struct Image {
    const float* rgb;
    int W, H;
};
static inline int wrapIndex(int i, int n) {
    int m = i % n;
    return m < 0 ? m + n : m;
}
static inline vec4 texelRGB(const Image& im, int x, int y) {
    const float* p = im.rgb + 3 * ((size_t)wrapIndex(y, im.H) * im.W + wrapIndex(x, im.W));
    return vec4(p[0], p[1], p[2], 1.0f);
}
static inline void linearCoord(real u, int size, int& i0, real& t) {
    real f = u * real((float)size) - real(0.5f);
    if (!(f.v == f.v) || !(f.v > -1e9f && f.v < 1e9f)) { i0 = 0; t = real(0.0f); return; }
    real fl = floor(f);
    i0 = (int)fl.v;
    t = floor((f - fl) * real(256.0f) + real(0.5f)) * real(1.0f / 256.0f);
}
static inline vec4 textureLinear(const Image& im, vec2 uv) {
    int x0, y0;
    real tx, ty;
    linearCoord(uv.x, im.W, x0, tx);
    linearCoord(uv.y, im.H, y0, ty);
    vec4 c00 = texelRGB(im, x0, y0), c10 = texelRGB(im, x0 + 1, y0);
    vec4 c01 = texelRGB(im, x0, y0 + 1), c11 = texelRGB(im, x0 + 1, y0 + 1);
    vec4 a = c00 * (real(1.0f) - tx) + c10 * tx;
    vec4 b = c01 * (real(1.0f) - tx) + c11 * tx;
    return a * (real(1.0f) - ty) + b * ty;
}

struct FilterParams {
    int id;
    float gamma_c;        // FILTER_GAMMA_C
    float radius[2];      // FILTER_WINDOW_RADIUS / FILTER_WAVELET_R
    float weights[16];    // windowWeightTable
};

// tonemapping.glsl:1-4
static inline vec3 tonemapping(vec3 color) {
    vec3 x = max(vec3(0.0f), color - real(0.004f));
    return (x * (real(6.2f) * x + real(0.5f))) / (x * (real(6.2f) * x + real(1.7f)) + real(0.06f));
}
// window.glsl:1-8
static inline vec3 windowSampler(const Image& colorMap, vec2 coord, int& count) {
    if (coord.x < real(0.0f) || coord.x > real(1.0f) || coord.y < real(0.0f) || coord.y > real(1.0f)) return vec3(0.0f, 0.0f, 0.0f);
    count++;
    return textureLinear(colorMap, coord).rgb();
}
// window.glsl:10-24
static inline vec3 window(const Image& colorMap, vec2 coord, real i, real j, int& count) {
    count = 0;
    vec2 x = vec2(i / real((float)colorMap.W), 0.0f);
    vec2 y = vec2(0.0f, j / real((float)colorMap.H));
    vec3 color = vec3(0.0f, 0.0f, 0.0f);
    color += windowSampler(colorMap, coord + x + y, count);
    color += windowSampler(colorMap, coord + x - y, count);
    color += windowSampler(colorMap, coord - x + y, count);
    color += windowSampler(colorMap, coord - x - y, count);
    return color;
}
// window.glsl:26-44.  Note the table lookup windowWeightTable[i*j+j] (sic).
static inline vec4 pixelFilterWindow(const Image& colorMap, const FilterParams& fp, vec2 texCoord) {
    vec3 color = vec3(0.0f, 0.0f, 0.0f);
    real weightSum = 0.0f;
    for (int i = 0; i < 4; i++) {
        for (int j = 0; j < 4; j++) {
            int count;
            vec3 tmpColor = window(colorMap, texCoord, (real((float)j) + real(0.5f)) * real(fp.radius[0]) / real(4.0f),
                                   (real((float)i) + real(0.5f)) * real(fp.radius[1]) / real(4.0f), count);
            real weight = real(fp.weights[i * j + j]);
            weightSum += weight * real((float)count);
            color += tmpColor * weight;
        }
    }
    return vec4(color / weightSum, 1.0f);
}
// wavelet.glsl:5-20.  `ntmp` is fetched but unused; the "normal" weight reuses the colour difference.
static inline real waveletW(const Image& colorMap, const Image& positionMap, vec2 uv, real stepwidth, real h, vec4 cval,
                            vec4 pval, vec4& ctmp) {
    ctmp = textureLinear(colorMap, uv);
    vec4 t = cval - ctmp;
    real dist2 = dot(t, t);
    real c_w = min(exp(-(dist2) / real(4.0f)), real(1.0f));
    dist2 = max(dot(t, t) / (stepwidth * stepwidth), real(0.0f));
    real n_w = min(exp(-(dist2) / real(128.0f)), real(1.0f));
    vec4 ptmp = textureLinear(positionMap, uv);
    t = pval - ptmp;
    dist2 = dot(t, t);
    real p_w = min(exp(-(dist2) / real(1.0f)), real(1.0f));
    real weight = c_w * n_w * p_w * h;
    ctmp *= weight;
    return weight;
}
// wavelet.glsl:22-54.  pow(2.0, float(n)) is taken as exactly 1, 2, 4.
static inline vec4 pixelFilterWavelet(const Image& colorMap, const Image& positionMap, const FilterParams& fp, vec2 texCoord) {
    vec4 color = vec4(0.0f);
    real weightSum = 0.0f;
    vec4 cval = textureLinear(colorMap, texCoord);
    vec4 pval = textureLinear(positionMap, texCoord);
    const float h[5] = {0.375f, 0.25f, 0.0625f, 0.0625f, 0.25f};
    real Rx = fp.radius[0], Ry = fp.radius[1];
    real Wf = (float)colorMap.W, Hf = (float)colorMap.H;
    for (int n = 0; n < 3; n++) {
        real stepwidth = real((float)(1 << n)) - real(1.0f);
        int count = 0;
        for (int i = 0; i < 5; i++) {
            for (int j = 0; j < 5; j++, count++) {
                int delt = abs(count - 12);
                real _h = 0.0f;
                int step = toInt(stepwidth) + 1;
                if (delt % step == 0) _h = real(h[(delt / step) % 5]);
                if (_h == real(0.0f)) continue;
                vec2 uv = texCoord - vec2(Rx / Wf, Ry / Hf) +
                          vec2((real((float)j) + real(0.5f)) * Rx / (real(2.5f) * Wf), (real((float)i) + real(0.5f)) * Ry / (real(2.5f) * Hf));
                vec4 ctmp;
                real weight = waveletW(colorMap, positionMap, uv, stepwidth, _h, cval, pval, ctmp);
                weightSum += weight;
                color += ctmp;
            }
        }
    }
    return color / weightSum;
}

// fsrender.glsl:4-6 with the configured filter plugin.
static inline vec4 pixelFilter(const Image& colorMap, const Image& normalMap, const Image& positionMap, const FilterParams& fp,
                               vec2 texCoord) {
    switch (fp.id) {
        case F_COLOR: return vec4(textureLinear(colorMap, texCoord).rgb(), 1.0f);
        case F_GAMMA: return vec4(pow(textureLinear(colorMap, texCoord).rgb(), vec3(real(1.0f) / real(fp.gamma_c))), 1.0f);
        case F_TONEMAPPING: return vec4(tonemapping(textureLinear(colorMap, texCoord).rgb()), 1.0f);
        case F_NORMAL: return vec4(textureLinear(normalMap, texCoord).rgb(), 1.0f);
        case F_POSITION: return vec4(textureLinear(positionMap, texCoord).rgb(), 1.0f);
        case F_WAVELET: return pixelFilterWavelet(colorMap, positionMap, fp, texCoord);
        case F_BOX: case F_GAUSSIAN: case F_MITCHELL: case F_SINC: case F_TRIANGLE:
            return pixelFilterWindow(colorMap, fp, texCoord);
    }
    return vec4(0.0f);
}

extern "C" int oracle_present(int W, int H, const float* color, const float* normal, const float* position, int filter_id,
                              float gamma_c, const float* radius, const float* weights, float* out_rgba, uint8_t* out_rgba8) {
    FilterParams fp;
    fp.id = filter_id;
    fp.gamma_c = gamma_c;
    fp.radius[0] = radius ? radius[0] : 0.0f;
    fp.radius[1] = radius ? radius[1] : 0.0f;
    for (int i = 0; i < 16; i++) fp.weights[i] = weights ? weights[i] : 0.0f;
    Image c{color, W, H}, n{normal, W, H}, p{position, W, H};
#pragma omp parallel for schedule(dynamic, 4)
    for (int y = 0; y < H; y++) {
        for (int x = 0; x < W; x++) {
            // vsrender.glsl:5 texCoord = vertex.xy*0.5+0.5, interpolated to the pixel centre.
            vec2 texCoord((real((float)x) + real(0.5f)) / real((float)W), (real((float)y) + real(0.5f)) / real((float)H));
            vec4 v = pixelFilter(c, n, p, fp, texCoord);
            size_t i = (size_t)y * W + x;
            const real comp[4] = {v.x, v.y, v.z, v.w};
            for (int k = 0; k < 4; k++) {
                if (out_rgba) out_rgba[4 * i + k] = comp[k].v;
                if (out_rgba8) {
                    float f = comp[k].v;
                    f = (f == f) ? dm_min(dm_max(f, 0.0f), 1.0f) : 0.0f;  // canvas: clamp, NaN -> 0
                    out_rgba8[4 * i + k] = (uint8_t)floorf(f * 255.0f + 0.5f);
                }
            }
        }
    }
    return 0;
}
#endif
