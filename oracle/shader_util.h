// oracle/shader_util.h -- TEST INFRASTRUCTURE ONLY.  Restates (reference paths under /root/reference):
//   src/shader/const/define.glsl:1-64, const/struct.glsl:1-18,
//   src/shader/util/random.glsl:1-18, util/sampler.glsl:1-54, util/texhelper.glsl:1-36,
//   util/utility.glsl:1-60, and the uniform block of src/core/shader.js:41-55.
#ifndef SAIL_ORACLE_SHADER_UTIL_H
#define SAIL_ORACLE_SHADER_UTIL_H
#include "glsl.h"

// ---- const/define.glsl:1-64 ------------------------------------------------------------------
#define OBJECTS_LENGTH 17.0f
#define LIGHTS_LENGTH 17.0f
#define TEX_PARAMS_LENGTH 15.0f
#define MAX_DISTANCE 1e5f
#define MAXBOUNCES 5
#define EPSILON 1e-5f
#define ONEMINUSEPSILON 0.9999f
#define INF 1e5f
#define PI 3.141592653589793f
#define INVPI 0.3183098861837907f
#define INV2PI 0.159154943091895f
#define INV4PI 0.079577471545947f
#define PIOVER2 1.570796326794896f
#define PIOVER4 0.785398163397448f
#define SQRT2 1.414213562373095f

enum { CUBE = 1, SPHERE = 2, RECTANGLE = 3, CONE = 4, CYLINDER = 5, DISK = 6, HYPERBOLOID = 7, PARABOLOID = 8, CORNELLBOX = 9 };
enum { AREA = 0, POINT = 1, SPOT = 2 };
enum { MATTE = 1, MIRROR = 2, METAL = 3, GLASS = 4 };
enum { UNIFORM_COLOR = 0, CHECKERBOARD = 5, CHECKERBOARD2 = 7, BILERP = 8, MIXF = 9, SCALE = 10, UVF = 11 };
enum { NOOP = 0, CONDUCTOR = 1, DIELECTRIC = 2 };
enum { BECKMANN = 1, TROWBRIDGEREITZ = 2 };

#define BLACK vec3(0.0f, 0.0f, 0.0f)
#define WHITE vec3(1.0f, 1.0f, 1.0f)
#define GREY vec3(0.5f, 0.5f, 0.5f)
#define RED vec3(0.75f, 0.25f, 0.25f)
#define BLUE vec3(0.25f, 0.25f, 0.75f)
#define GREEN vec3(0.25f, 0.75f, 0.25f)
#define OBJECT_SPACE_N vec3(0.0f, 1.0f, 0.0f)
#define OBJECT_SPACE_S vec3(0.0f, 0.0f, -1.0f)
#define OBJECT_SPACE_T vec3(1.0f, 0.0f, 0.0f)

// ---- const/struct.glsl:1-18 ------------------------------------------------------------------
struct Intersect {
    real d;
    vec3 hit;
    vec3 normal;
    vec3 dpdu, dpdv;
    bool into = false;
    real matIndex;
    vec3 sc;
    vec3 emission;
    real seed;
    int index = 0;
    int matCategory = 0;
};
struct Ray {
    vec3 origin;
    vec3 dir;
    Ray() {}
    Ray(vec3 o, vec3 d) : origin(o), dir(d) {}
};

// ---- uniforms + samplers (src/core/shader.js:41-55) and gl_FragCoord --------------------------
struct Uniforms {
    int n = 0, ln = 0, tn = 0;
    real textureWeight, timeSinceStart_;
    mat4 matrix;
    vec3 eye;
    sampler2D objects_{nullptr, 18, 0}, texParams_{nullptr, 16, 0}, lights_{nullptr, 18, 0};
    int maxBounces = MAXBOUNCES;  // define.glsl:6; config 5 of BASELINE.json raises it to 16
    int rrStart = 0;              // Russian roulette from this bounce on (0 = off, the reference's behaviour): see trace()
};
static Uniforms U;
static thread_local vec3 gl_FragCoord;
#define n_ (U.n)
#define ln_ (U.ln)
#define tn_ (U.tn)
#define objects (U.objects_)
#define texParams (U.texParams_)
#define lights (U.lights_)
#define timeSinceStart (U.timeSinceStart_)

// Row coordinate helpers: `float(i)/float(n-1)` (shader.shape.js:34,44,56; shader.light.js:20) and
// `value/float(tn-1)` (cube.glsl:19-20 ...).  Pure table addressing: not counted as arithmetic.
#ifdef SAIL_ORACLE_IEEE_DIV
static inline real rowCoord(int i, int count) { return real((float)i / (float)(count - 1)); }
static inline real rowCoordF(real v, int count) { return real(v.v / (float)(count - 1)); }
#else
static inline real rowCoord(int i, int count) { return real(dm_div((float)i, (float)(count - 1))); }
static inline real rowCoordF(real v, int count) { return real(dm_div(v.v, (float)(count - 1))); }
#endif

// ---- util/texhelper.glsl:1-36 ----------------------------------------------------------------
static inline vec2 convert(vec2 pos, float width) {
    pos.x = real(pos.x.v / width);
    return pos;
}
static inline int readInt(const sampler2D& tex, vec2 pos, float width) { return toInt(texture(tex, convert(pos, width))); }
static inline real readFloat(const sampler2D& tex, vec2 pos, float width) { return texture(tex, convert(pos, width)); }
static inline bool readBool(const sampler2D& tex, vec2 pos, float width) { return readInt(tex, pos, width) == 1; }
static inline vec2 readVec2(const sampler2D& tex, vec2 pos, float width) {
    vec2 result;
    pos = convert(pos, width);
    result.x = texture(tex, pos);
    pos.x = real(pos.x.v + 1.0f / width);
    result.y = texture(tex, pos);
    return result;
}
static inline vec3 readVec3(const sampler2D& tex, vec2 pos, float width) {
    vec3 result;
    pos = convert(pos, width);
    result.x = texture(tex, pos);
    pos.x = real(pos.x.v + 1.0f / width);
    result.y = texture(tex, pos);
    pos.x = real(pos.x.v + 1.0f / width);
    result.z = texture(tex, pos);
    return result;
}

// ---- util/random.glsl:1-18 -------------------------------------------------------------------
static inline real random(vec3 scale, real seed) {
    return fract(sin(dot(gl_FragCoord + seed, scale)) * real(43758.5453f) + seed);
}
static inline vec2 random2(real seed) {
    return vec2(fract(sin(dot(gl_FragCoord + seed, vec3(12.9898f, 78.233f, 151.7182f))) * real(43758.5453f) + seed),
                fract(sin(dot(gl_FragCoord + seed, vec3(63.7264f, 10.873f, 623.6736f))) * real(43758.5453f) + seed));
}
static inline int randomInt(real seed, int min_, int max_) {
    return min_ + toInt(fract(sin(dot(gl_FragCoord + seed, vec3(12.9898f, 78.233f, 151.7182f))) * real(43758.5453f) + seed) *
                        real((float)(max_ - min_)));
}

// ---- util/sampler.glsl:1-54 (cosineSampleHemisphere2, uniformSampleDisk/Cone/Triangle are unused) ---
static inline vec3 uniformSampleSphere(vec2 u) {
    real z = real(1.0f) - real(2.0f) * u.x;
    real r = sqrt(real(1.0f) - z * z);
    real angle = real(2.0f) * real(PI) * u.y;
    return vec3(r * cos(angle), r * sin(angle), z);
}
static inline vec3 cosineSampleHemisphere(vec2 u) {
    real r = sqrt(u.x);
    real angle = real(2.0f) * real(PI) * u.y;
    return vec3(r * cos(angle), r * sin(angle), sqrt(real(1.0f) - u.x));
}
static inline vec2 concentricSampleDisk(vec2 u) {
    real uOffset = real(2.0f) * u.x - real(1.0f);
    real vOffset = real(2.0f) * u.y - real(1.0f);
    if (uOffset == real(0.0f) && vOffset == real(0.0f)) return vec2(0.0f, 0.0f);
    real theta, r;
    if (abs(uOffset) > abs(vOffset)) {
        r = uOffset;
        theta = (vOffset / uOffset) * real(PIOVER4);
    } else {
        r = vOffset;
        theta = real(PIOVER2) - (uOffset / vOffset) * real(PIOVER4);
    }
    return r * vec2(cos(theta), sin(theta));
}

// ---- util/utility.glsl:1-60 ------------------------------------------------------------------
static inline vec3 worldToLocal(vec3 v, vec3 ns, vec3 ss, vec3 ts) { return vec3(dot(v, ss), dot(v, ts), dot(v, ns)); }
static inline vec3 localToWorld(vec3 v, vec3 ns, vec3 ss, vec3 ts) {
    return vec3(ss.x * v.x + ts.x * v.y + ns.x * v.z, ss.y * v.x + ts.y * v.y + ns.y * v.z,
                ss.z * v.x + ts.z * v.y + ns.z * v.z);
}
static inline vec3 ensure3byW(vec4 v) { return vec3(v.x / v.w, v.y / v.w, v.z / v.w); }
static inline void swap(real& f1, real& f2) {
    real tmp = f1;
    f1 = f2;
    f2 = tmp;
}
static inline bool quadratic(real A, real B, real C, real& t0, real& t1) {
    real discrim = B * B - real(4.0f) * A * C;
    if (discrim < real(0.0f)) return false;
    real rootDiscrim = sqrt(discrim);
    real q;
    if (B < real(0.0f))
        q = real(-0.5f) * (B - rootDiscrim);
    else
        q = real(-0.5f) * (B + rootDiscrim);
    t0 = q / A;
    t1 = C / q;
    if (t0 > t1) swap(t0, t1);
    return true;
}
static inline vec3 sphericalDirection(real sinTheta, real cosTheta, real phi) {
    return vec3(sinTheta * cos(phi), sinTheta * sin(phi), cosTheta);
}
static inline bool equalZero(real x) { return x < real(1e-3f) && x > real(-1e-3f); }

#endif
