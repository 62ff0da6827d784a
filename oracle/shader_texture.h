// oracle/shader_texture.h -- TEST INFRASTRUCTURE ONLY.  Restates src/shader/texture/:
//   shader.texture.js:22-29 (getSurfaceColor dispatcher), checkerboard.glsl:1-20,
//   checkerboard2.glsl:1-16, bilerp.glsl:1-14, mixf.glsl:1-13, scale.glsl:1-11, uvf.glsl:1-3.
// noise.glsl is never called by the reference and is not restated.
#ifndef SAIL_ORACLE_SHADER_TEXTURE_H
#define SAIL_ORACLE_SHADER_TEXTURE_H
#include "shader_util.h"

// checkerboard.glsl:1-20
static inline vec3 checkerboard(vec3 hit, vec2 uv, real texIndex) {
    real size = readFloat(texParams, vec2(1.0f, texIndex), TEX_PARAMS_LENGTH);
    real lineWidth = readFloat(texParams, vec2(2.0f, texIndex), TEX_PARAMS_LENGTH);
    real width = real(0.5f) * lineWidth / size;
    real fx = uv.x / size - floor(uv.x / size), fy = uv.y / size - floor(uv.y / size);
    bool in_outline = (fx < width || fx > real(1.0f) - width) || (fy < width || fy > real(1.0f) - width);
    if (!in_outline) return WHITE;
    return GREY;
}
// checkerboard2.glsl:1-16
static inline vec3 checkerboard2(vec3 hit, vec2 uv, real texIndex) {
    vec3 color1 = readVec3(texParams, vec2(1.0f, texIndex), TEX_PARAMS_LENGTH);
    vec3 color2 = readVec3(texParams, vec2(4.0f, texIndex), TEX_PARAMS_LENGTH);
    real size = readFloat(texParams, vec2(7.0f, texIndex), TEX_PARAMS_LENGTH);
    uv = vec2(floor(uv.x / size), floor(uv.y / size));
    if (toInt(uv.x + uv.y) % 2 == 0) return color1;
    return color2;
}
// bilerp.glsl:1-14.  Does not compile in the reference (`color11` undeclared):
// no reference behaviour exists; the evident intent (bilinear blend of 4 constants) is implemented.
static inline vec3 bilerp(vec3 hit, vec2 uv, real texIndex) {
    vec3 color00 = readVec3(texParams, vec2(1.0f, texIndex), TEX_PARAMS_LENGTH);
    vec3 color01 = readVec3(texParams, vec2(4.0f, texIndex), TEX_PARAMS_LENGTH);
    vec3 color10 = readVec3(texParams, vec2(7.0f, texIndex), TEX_PARAMS_LENGTH);
    vec3 color11 = readVec3(texParams, vec2(10.0f, texIndex), TEX_PARAMS_LENGTH);
    return (real(1.0f) - uv.x) * (real(1.0f) - uv.y) * color00 + (real(1.0f) - uv.x) * (uv.y) * color01 +
           (uv.x) * (real(1.0f) - uv.y) * color10 + (uv.x) * (uv.y) * color11;
}
// mixf.glsl:1-13
static inline vec3 mixf(vec3 hit, vec2 uv, real texIndex) {
    vec3 color1 = readVec3(texParams, vec2(1.0f, texIndex), TEX_PARAMS_LENGTH);
    vec3 color2 = readVec3(texParams, vec2(4.0f, texIndex), TEX_PARAMS_LENGTH);
    real amount = readFloat(texParams, vec2(7.0f, texIndex), TEX_PARAMS_LENGTH);
    return (real(1.0f) - amount) * color1 + amount * color2;
}
// scale.glsl:1-11
static inline vec3 scale(vec3 hit, vec2 uv, real texIndex) {
    vec3 color1 = readVec3(texParams, vec2(1.0f, texIndex), TEX_PARAMS_LENGTH);
    vec3 color2 = readVec3(texParams, vec2(4.0f, texIndex), TEX_PARAMS_LENGTH);
    return color1 * color2;
}
// uvf.glsl:1-3
static inline vec3 uvf(vec3 hit, vec2 uv, real texIndex) { return vec3(uv.x - floor(uv.x), uv.y - floor(uv.y), 0.0f); }

// shader.texture.js:22-29
static inline vec3 getSurfaceColor(vec3 hit, vec2 uv, real texIndex) {
    int texCategory = readInt(texParams, vec2(0.0f, texIndex), TEX_PARAMS_LENGTH);
    if (texCategory == UNIFORM_COLOR) return readVec3(texParams, vec2(1.0f, texIndex), TEX_PARAMS_LENGTH);
    else if (texCategory == CHECKERBOARD) return checkerboard(hit, uv, texIndex);
    else if (texCategory == CHECKERBOARD2) return checkerboard2(hit, uv, texIndex);
    else if (texCategory == BILERP) return bilerp(hit, uv, texIndex);
    else if (texCategory == MIXF) return mixf(hit, uv, texIndex);
    else if (texCategory == SCALE) return scale(hit, uv, texIndex);
    else if (texCategory == UVF) return uvf(hit, uv, texIndex);
    return BLACK;
}
#endif
