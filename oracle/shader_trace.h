// oracle/shader_trace.h -- TEST INFRASTRUCTURE ONLY.  Restates src/shader/light/shader.light.js:12-31,
// light/area.glsl:1-23, light/point.glsl:1-20, light/spot.glsl:1-39, trace/path.glsl:1-38,
// main/vstrace.glsl:4-7 and main/fstrace.glsl:6-17.
#ifndef SAIL_ORACLE_SHADER_TRACE_H
#define SAIL_ORACLE_SHADER_TRACE_H
#include "shader_material.h"
#include "shader_shape.h"

// ---- shader.light.js:24-31 -------------------------------------------------------------------
static inline bool testShadow(const Ray& ray) {
    Intersect ins = intersectObjects(ray);
    if (ins.d > real(EPSILON) && ins.d < real(ONEMINUSEPSILON)) return true;
    return false;
}

// ---- area.glsl:1-23 --------------------------------------------------------------------------
struct Area {
    vec3 emission;
    int index = 0;
};
static inline Area parseArea(real index) {
    Area area;
    area.index = readInt(lights, vec2(1.0f, index), LIGHTS_LENGTH);
    area.emission = readVec3(lights, vec2(2.0f, index), LIGHTS_LENGTH);
    return area;
}
static inline vec3 area_sample(const Area& area, real seed, vec3 hit, vec3 insNormal) {
    vec3 normal, p;
    real pdf;
    p = sampleGeometry(random2(seed), area.index, normal, pdf);
    vec3 toLight = p - hit;
    vec3 normToLight = normalize(toLight);
    if (testShadow(Ray(hit, toLight))) return BLACK;
    return area.emission * max(real(0.0f), dot(normal, -normToLight)) * max(real(0.0f), dot(normToLight, insNormal)) / pdf;
}

// ---- point.glsl:1-20 -------------------------------------------------------------------------
struct Point {
    vec3 from, emission;
};
static inline Point parsePoint(real index) {
    Point point;
    point.from = readVec3(lights, vec2(1.0f, index), LIGHTS_LENGTH);
    point.emission = readVec3(lights, vec2(4.0f, index), LIGHTS_LENGTH);
    return point;
}
static inline vec3 point_sample(const Point& point, real seed, vec3 hit, vec3 insNormal) {
    vec3 p = point.from + uniformSampleSphere(random2(seed)) * real(0.1f);
    vec3 toLight = p - hit;
    if (testShadow(Ray(hit, toLight))) return BLACK;
    return point.emission * max(real(0.0f), dot(normalize(toLight), insNormal));
}

// ---- spot.glsl:1-39 --------------------------------------------------------------------------
struct Spot {
    vec3 emission;
    real cosTotalWidth, cosFalloffStart;
    vec3 from;
};
static inline Spot parseSpot(real index) {
    Spot spot;
    spot.cosTotalWidth = readFloat(lights, vec2(1.0f, index), LIGHTS_LENGTH);
    spot.cosFalloffStart = readFloat(lights, vec2(2.0f, index), LIGHTS_LENGTH);
    spot.from = readVec3(lights, vec2(3.0f, index), LIGHTS_LENGTH);
    spot.emission = readVec3(lights, vec2(6.0f, index), LIGHTS_LENGTH);
    return spot;
}
static inline real falloff(const Spot& spot, vec3 w) {
    vec3 wl = w;
    real cosTheta_ = -wl.y;
    if (cosTheta_ < spot.cosTotalWidth) return real(0.0f);
    if (cosTheta_ >= spot.cosFalloffStart) return real(1.0f);
    real delta = (cosTheta_ - spot.cosTotalWidth) / (spot.cosFalloffStart - spot.cosTotalWidth);
    real delta2 = delta * delta;
    return delta2 * delta2;
}
static inline vec3 spot_sample(const Spot& spot, real seed, vec3 hit, vec3 insNormal) {
    vec3 p = spot.from;
    vec3 toLight = p - hit;
    if (testShadow(Ray(hit, toLight))) return BLACK;
    vec3 normToLight = normalize(toLight);
    real d = length(toLight);
    return spot.emission * falloff(spot, -normToLight) * max(real(0.0f), dot(normalize(toLight), insNormal)) / (d * d);
}

// ---- shader.light.js:12-22 -------------------------------------------------------------------
static inline vec3 light_sample(const Intersect& ins) {
    vec3 fpdf = BLACK;
    // ln == 0: no light plugin is compiled in and the lights texture has zero height, so no arm can
    // fire (scene.js:104-110, tracer.js:76-80).
    if (ln_ == 0) return fpdf;
    int index = randomInt(ins.seed, 0, ln_);
    // shader.light.js:15 passes the raw int as a normalised row coordinate: index 0 -> row 0, any
    // index >= 1 -> clamped to the last row.  The texture emulation reproduces it.
    int lightCategory = readInt(lights, vec2(0.0f, real((float)index)), TEX_PARAMS_LENGTH);
    if (false) {
    } else if (lightCategory == AREA) {
        Area area = parseArea(rowCoord(index, ln_));
        fpdf = area_sample(area, ins.seed, ins.hit, ins.normal);
    } else if (lightCategory == POINT) {
        Point point = parsePoint(rowCoord(index, ln_));
        fpdf = point_sample(point, ins.seed, ins.hit, ins.normal);
    } else if (lightCategory == SPOT) {
        Spot spot = parseSpot(rowCoord(index, ln_));
        fpdf = spot_sample(spot, ins.seed, ins.hit, ins.normal);
    }
    return fpdf;
}

// ---- path.glsl:1-14 --------------------------------------------------------------------------
static inline vec3 shade(const Intersect& ins, vec3 wo, vec3& wi, vec3& fpdf) {
    vec3 f, direct = BLACK;
    vec3 ss = normalize(ins.dpdu), ts = cross(ins.normal, ss);
    wo = worldToLocal(wo, ins.normal, ss, ts);
    fpdf = clamp(material(ins, wo, wi, f), BLACK, WHITE);
    wi = localToWorld(wi, ins.normal, ss, ts);
    if (ins.emission == BLACK && ins.matCategory == MATTE) direct += light_sample(ins) * f;
    return ins.emission + direct;
}

// Optional per-bounce dump used by the replay parity tests (hit id / t / ray of every bounce).
struct BounceDump {
    int* hitId;     // [maxBounces]  -1 = miss / not reached
    float* t;       // [maxBounces]
    float* ray;     // [maxBounces*6] origin, dir of the ray traced at that bounce
};

// ---- path.glsl:16-38 -------------------------------------------------------------------------
static inline void trace(Ray ray, int maxDeepth, vec3& e, vec3& n, vec3& p, int* firstHit, real* firstT, BounceDump* dump) {
    vec3 fpdf = WHITE;
    e = BLACK;
    int deepth = 0;
    while (deepth++ < maxDeepth) {
        Intersect ins = intersectObjects(ray);
        ins.seed = timeSinceStart + real((float)deepth);
        if (dump) {
            dump->hitId[deepth - 1] = (ins.d >= real(MAX_DISTANCE)) ? -1 : ins.index;
            dump->t[deepth - 1] = ins.d.v;
            float* r = dump->ray + 6 * (deepth - 1);
            r[0] = ray.origin.x.v; r[1] = ray.origin.y.v; r[2] = ray.origin.z.v;
            r[3] = ray.dir.x.v; r[4] = ray.dir.y.v; r[5] = ray.dir.z.v;
        }
        if (deepth == 1) {
            *firstHit = (ins.d >= real(MAX_DISTANCE)) ? -1 : ins.index;
            *firstT = ins.d;
        }
        if (ins.d >= real(MAX_DISTANCE)) break;
        if (deepth == 1) {
            n = ins.normal;
            p = ins.hit;
        }
        vec3 wi;
        vec3 _fpdf;
        e += shade(ins, -ray.dir, wi, _fpdf) * fpdf;
        fpdf *= _fpdf;
        // Russian roulette -- NOT in the reference (fixed MAXBOUNCES, define.glsl:6); an opt-in extension the build's brief names
        // (BASELINE.json north_star), off unless rrStart > 0.  From bounce rrStart on a path survives with probability
        // q = clamp(max(fpdf), 0.05, 1) and its throughput is divided by q; the decision draws one more number from the
        // shader's own hash (random(), util/random.glsl:1-4) with a third scale vector.
        if (U.rrStart > 0 && deepth >= U.rrStart && deepth < maxDeepth) {
            real q = clamp(max(fpdf.x, max(fpdf.y, fpdf.z)), real(0.05f), real(1.0f));
            real u = random(vec3(37.1234f, 91.4567f, 231.9876f), ins.seed);
            if (!(u < q)) break;
            fpdf = fpdf * (real(1.0f) / q);
        }
        real outdot = dot(ins.normal, wi);
        ray.origin = ins.hit + ins.normal * (outdot > real(EPSILON) ? real(0.0001f) : real(-0.0001f));
        ray.dir = wi;
    }
}

// ---- vstrace.glsl:4-7 + the rasteriser's interpolation of `raydir` -------------
// The quad is the 4-vertex strip (-1,-1),(-1,+1),(+1,-1),(+1,+1) (tracer.js:16-21): triangles
// (v0,v1,v2) and (v2,v1,v3) share the diagonal v1-v2.  A varying is interpolated linearly (w = 1 at every
// vertex) within the triangle that contains the pixel centre.  Interpolation is fixed-function
// hardware in the reference; it is defined here with a = (x+.5)/W, b = (y+.5)/H as
//   a + b < 1 : d0 + a*(d2-d0) + b*(d1-d0)        (lower-left triangle)
//   else      : d3 + (1-a)*(d1-d3) + (1-b)*(d2-d3)  (upper-right triangle)
struct FrameRays {
    vec3 d0, d1, d2, d3;
};
static inline vec3 cornerRay(real vx, real vy) {
    vec4 gl_Position = vec4(vx, vy, 0.0f, 1.0f);
    return normalize(ensure3byW(U.matrix * gl_Position) - U.eye);
}
static inline FrameRays makeFrameRays() {
    FrameRays fr;
    fr.d0 = cornerRay(-1.0f, -1.0f);
    fr.d1 = cornerRay(-1.0f, 1.0f);
    fr.d2 = cornerRay(1.0f, -1.0f);
    fr.d3 = cornerRay(1.0f, 1.0f);
    return fr;
}
static inline vec3 interpolateRay(const FrameRays& fr, int x, int y, int W, int H) {
    real a = (real((float)x) + real(0.5f)) / real((float)W);
    real b = (real((float)y) + real(0.5f)) / real((float)H);
    if (a + b < real(1.0f)) return fr.d0 + a * (fr.d2 - fr.d0) + b * (fr.d1 - fr.d0);
    return fr.d3 + (real(1.0f) - a) * (fr.d1 - fr.d3) + (real(1.0f) - b) * (fr.d2 - fr.d3);
}

// ---- fstrace.glsl:6-17 -----------------------------------------------------------------------
struct PixelOut {
    vec3 color, normal, position;
    int hitId;
    real depth;
};
// NaN in the per-sample radiance is mapped to 0 before accumulation.
static inline real nanToZero(real v) { return (v.v == v.v) ? v : real(0.0f); }
static inline PixelOut fstrace_main(const FrameRays& fr, int x, int y, int W, int H, vec3 cache, BounceDump* dump) {
    gl_FragCoord = vec3((float)x + 0.5f, (float)y + 0.5f, 0.5f);
    vec3 e, n, p;
    Ray ray(U.eye, interpolateRay(fr, x, y, W, H));
    PixelOut out;
    trace(ray, U.maxBounces, e, n, p, &out.hitId, &out.depth, dump);
    e = vec3(nanToZero(e.x), nanToZero(e.y), nanToZero(e.z));
    out.color = mix(e, cache, U.textureWeight);
    out.normal = n / real(2.0f) + real(0.5f);
    out.position = normalize(p);
    return out;
}
#endif
