// oracle/glsl.h -- TEST INFRASTRUCTURE ONLY (the CPU oracle; never linked into the product).
//
// A tiny GLSL ES 3.00 value model in C++ so that the reference shaders under
// /root/reference/src/shader/** can be restated function by function (oracle/shader_*.h).
//
// PARITY UNPINNED: the reference has no tests, golden vectors or runnable CPU path for its shaders
// (SURVEY.md 4, 8(c)); GLSL leaves the precision of '/', sqrt and every transcendental implementation
// defined.  This file therefore DEFINES the arithmetic the oracle (and the CUDA kernels) use:
//   * `real` is IEEE-754 binary32, round-to-nearest, no contraction (build with -ffp-contract=off).
//   * a / b      := a * rcp(b), rcp correctly rounded  (what shader compilers emit; GLSL ES 3.00 4.5.1).
//   * sqrt       := correctly rounded.
//   * sin, cos, tan, atan, acos, asin, exp, log, pow := sail_b200/csrc/detmath.h (shared, deterministic).
//   * min/max    := a NaN operand yields the other operand; -0 orders below +0 (IEEE 754-2019 minimumNumber /
//                   maximumNumber, which is what the sm_100a FMNMX instruction computes).
//   * normalize(v) := v * rcp(sqrt(dot(v,v))); length(v) := sqrt(dot(v,v)).
//   * unwritten `out` parameters and uninitialised locals read as 0: `real` and the
//     vector types zero-initialise.
// With -DSAIL_COUNT_OPS every arithmetic operation on `real` bumps a thread-local counter
// (SURVEY.md 8(d) "OPS_alg": + - * / min max abs floor compare sqrt and each transcendental = 1).
#ifndef SAIL_ORACLE_GLSL_H
#define SAIL_ORACLE_GLSL_H

#include <stdint.h>
#include "../sail_b200/csrc/detmath.h"

#ifdef SAIL_COUNT_OPS
extern thread_local uint64_t g_sail_ops;
#define SAIL_OP(n) (g_sail_ops += (n))
#else
#define SAIL_OP(n) ((void)0)
#endif

struct real {
    float v;
    real() : v(0.0f) {}
    real(float f) : v(f) {}
    real(double d) : v((float)d) {}
    real(int i) : v((float)i) {}
};

static inline real operator+(real a, real b) { SAIL_OP(1); return real(a.v + b.v); }
static inline real operator-(real a, real b) { SAIL_OP(1); return real(a.v - b.v); }
static inline real operator*(real a, real b) { SAIL_OP(1); return real(a.v * b.v); }
#ifdef SAIL_ORACLE_IEEE_DIV
// The other defensible reading of GLSL's '/' (SURVEY.md 8(c) first wrote "correctly rounded"): IEEE-754 division, and
// normalize(v) = v / length(v).  Built as liboracle_ieee.so ONLY to measure how far the two arithmetic models are apart
// (tests/test_cpu_model_sensitivity.py); parity tests use the a * rcp(b) model above.
static inline real operator/(real a, real b) { SAIL_OP(1); return real(a.v / b.v); }
#else
static inline real operator/(real a, real b) { SAIL_OP(1); return real(dm_div(a.v, b.v)); }
#endif
static inline real operator-(real a) { return real(-a.v); }
static inline real& operator+=(real& a, real b) { a = a + b; return a; }
static inline real& operator-=(real& a, real b) { a = a - b; return a; }
static inline real& operator*=(real& a, real b) { a = a * b; return a; }
static inline bool operator<(real a, real b) { SAIL_OP(1); return a.v < b.v; }
static inline bool operator>(real a, real b) { SAIL_OP(1); return a.v > b.v; }
static inline bool operator<=(real a, real b) { SAIL_OP(1); return a.v <= b.v; }
static inline bool operator>=(real a, real b) { SAIL_OP(1); return a.v >= b.v; }
static inline bool operator==(real a, real b) { SAIL_OP(1); return a.v == b.v; }
static inline bool operator!=(real a, real b) { SAIL_OP(1); return a.v != b.v; }

static inline real sqrt(real a) { SAIL_OP(1); return real(dm_sqrt(a.v)); }
static inline real abs(real a) { SAIL_OP(1); return real(dm_abs(a.v)); }
static inline real floor(real a) { SAIL_OP(1); return real(dm_floor(a.v)); }
static inline real min(real a, real b) { SAIL_OP(1); return real(dm_min(a.v, b.v)); }
static inline real max(real a, real b) { SAIL_OP(1); return real(dm_max(a.v, b.v)); }
static inline real clamp(real x, real lo, real hi) { return min(max(x, lo), hi); }
static inline real fract(real a) { return a - floor(a); }
static inline real mix(real x, real y, real a) { return x * (real(1.0f) - a) + y * a; }
static inline real sin(real a) { SAIL_OP(1); return real(dm_sin(a.v)); }
static inline real cos(real a) { SAIL_OP(1); return real(dm_cos(a.v)); }
static inline real tan(real a) { SAIL_OP(1); return real(dm_tan(a.v)); }
static inline real atan(real a) { SAIL_OP(1); return real(dm_atan(a.v)); }
static inline real atan(real y, real x) { SAIL_OP(1); return real(dm_atan2(y.v, x.v)); }
static inline real acos(real a) { SAIL_OP(1); return real(dm_acos(a.v)); }
static inline real asin(real a) { SAIL_OP(1); return real(dm_asin(a.v)); }
static inline real exp(real a) { SAIL_OP(1); return real(dm_exp(a.v)); }
static inline real log(real a) { SAIL_OP(1); return real(dm_log(a.v)); }
static inline real pow(real a, real b) { SAIL_OP(1); return real(dm_pow(a.v, b.v)); }
// GLSL int(float): truncation toward zero (undefined for NaN/out of range; defined here as 0).
static inline int toInt(real a) {
    if (!(a.v > -2147483648.0f && a.v < 2147483648.0f)) return 0;
    return (int)a.v;
}

struct vec2 {
    real x, y;
    vec2() {}
    vec2(real a, real b) : x(a), y(b) {}
    explicit vec2(real a) : x(a), y(a) {}
};
struct vec3 {
    real x, y, z;
    vec3() {}
    vec3(real a, real b, real c) : x(a), y(b), z(c) {}
    explicit vec3(real a) : x(a), y(a), z(a) {}
    vec3(real a, vec2 bc) : x(a), y(bc.x), z(bc.y) {}
    vec2 xy() const { return vec2(x, y); }
    vec2 yz() const { return vec2(y, z); }
    vec2 xz() const { return vec2(x, z); }
};
struct vec4 {
    real x, y, z, w;
    vec4() {}
    vec4(real a, real b, real c, real d) : x(a), y(b), z(c), w(d) {}
    explicit vec4(real a) : x(a), y(a), z(a), w(a) {}
    vec4(vec3 v, real d) : x(v.x), y(v.y), z(v.z), w(d) {}
    vec3 rgb() const { return vec3(x, y, z); }
};

static inline vec2 operator+(vec2 a, vec2 b) { return vec2(a.x + b.x, a.y + b.y); }
static inline vec2 operator-(vec2 a, vec2 b) { return vec2(a.x - b.x, a.y - b.y); }
static inline vec2 operator*(vec2 a, real s) { return vec2(a.x * s, a.y * s); }
static inline vec2 operator*(real s, vec2 a) { return vec2(s * a.x, s * a.y); }
static inline vec2 operator/(vec2 a, vec2 b) { return vec2(a.x / b.x, a.y / b.y); }

static inline vec3 operator+(vec3 a, vec3 b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline vec3 operator-(vec3 a, vec3 b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline vec3 operator*(vec3 a, vec3 b) { return vec3(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline vec3 operator/(vec3 a, vec3 b) { return vec3(a.x / b.x, a.y / b.y, a.z / b.z); }
static inline vec3 operator+(vec3 a, real s) { return vec3(a.x + s, a.y + s, a.z + s); }
static inline vec3 operator-(vec3 a, real s) { return vec3(a.x - s, a.y - s, a.z - s); }
static inline vec3 operator*(vec3 a, real s) { return vec3(a.x * s, a.y * s, a.z * s); }
static inline vec3 operator*(real s, vec3 a) { return vec3(s * a.x, s * a.y, s * a.z); }
static inline vec3 operator/(vec3 a, real s) { return vec3(a.x / s, a.y / s, a.z / s); }
static inline vec3 operator-(vec3 a) { return vec3(-a.x, -a.y, -a.z); }
static inline vec3& operator+=(vec3& a, vec3 b) { a = a + b; return a; }
static inline vec3& operator*=(vec3& a, vec3 b) { a = a * b; return a; }
static inline bool operator==(vec3 a, vec3 b) { return (a.x == b.x) && (a.y == b.y) && (a.z == b.z); }

static inline vec4 operator+(vec4 a, vec4 b) { return vec4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
static inline vec4 operator-(vec4 a, vec4 b) { return vec4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }
static inline vec4 operator*(vec4 a, real s) { return vec4(a.x * s, a.y * s, a.z * s, a.w * s); }
static inline vec4 operator/(vec4 a, real s) { return vec4(a.x / s, a.y / s, a.z / s, a.w / s); }
static inline vec4& operator+=(vec4& a, vec4 b) { a = a + b; return a; }
static inline vec4& operator*=(vec4& a, real s) { a = a * s; return a; }

// GLSL ES 3.00 8.4 geometric functions, evaluated left to right as written in the spec.
static inline real dot(vec3 a, vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline real dot(vec4 a, vec4 b) { return a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w; }
static inline vec3 cross(vec3 x, vec3 y) {
    return vec3(x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y);
}
static inline real length(vec3 a) { return sqrt(dot(a, a)); }
#ifdef SAIL_ORACLE_IEEE_DIV
static inline vec3 normalize(vec3 a) { return a / sqrt(dot(a, a)); }
#else
static inline vec3 normalize(vec3 a) { return a * (real(1.0f) / sqrt(dot(a, a))); }
#endif
static inline vec3 reflect(vec3 I, vec3 N) { return I - (real(2.0f) * dot(N, I)) * N; }
static inline vec3 refract(vec3 I, vec3 N, real eta) {
    real d = dot(N, I);
    real k = real(1.0f) - eta * eta * (real(1.0f) - d * d);
    if (k < real(0.0f)) return vec3(0.0f, 0.0f, 0.0f);
    return eta * I - (eta * d + sqrt(k)) * N;
}
static inline vec3 min(vec3 a, vec3 b) { return vec3(min(a.x, b.x), min(a.y, b.y), min(a.z, b.z)); }
static inline vec3 max(vec3 a, vec3 b) { return vec3(max(a.x, b.x), max(a.y, b.y), max(a.z, b.z)); }
static inline vec3 clamp(vec3 x, vec3 lo, vec3 hi) { return min(max(x, lo), hi); }
static inline vec3 sqrt(vec3 a) { return vec3(sqrt(a.x), sqrt(a.y), sqrt(a.z)); }
static inline vec3 pow(vec3 a, vec3 b) { return vec3(pow(a.x, b.x), pow(a.y, b.y), pow(a.z, b.z)); }
static inline vec3 mix(vec3 x, vec3 y, real a) { return x * (real(1.0f) - a) + y * a; }

// mat4 stored column-major like gl.uniformMatrix4fv(false, ...) (webgl.js:103, matrix.js:612-622).
struct mat4 {
    real m[16];
};
static inline vec4 operator*(const mat4& M, vec4 v) {
    return vec4(M.m[0] * v.x + M.m[4] * v.y + M.m[8] * v.z + M.m[12] * v.w,
                M.m[1] * v.x + M.m[5] * v.y + M.m[9] * v.z + M.m[13] * v.w,
                M.m[2] * v.x + M.m[6] * v.y + M.m[10] * v.z + M.m[14] * v.w,
                M.m[3] * v.x + M.m[7] * v.y + M.m[11] * v.z + M.m[15] * v.w);
}

// sampler2D over an R32F table, NEAREST + CLAMP_TO_EDGE (webgl.js:146-158).  Addressing is not
// counted as arithmetic.  A NaN coordinate resolves to texel 0.
struct sampler2D {
    const float* data;
    int width, height;
};
static inline int sail_texel(float coord, int size) {
    if (!(coord == coord)) return 0;
    float f = floorf(coord * (float)size);
    if (f < 0.0f) return 0;
    if (f > (float)(size - 1)) return size - 1;
    return (int)f;
}
static inline real texture(const sampler2D& t, vec2 pos) {
    if (t.width <= 0 || t.height <= 0) return real(0.0f);  // incomplete texture samples as 0
    int x = sail_texel(pos.x.v, t.width), y = sail_texel(pos.y.v, t.height);
    return real(t.data[y * t.width + x]);
}

#endif
