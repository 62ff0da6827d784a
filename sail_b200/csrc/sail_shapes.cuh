// sail_shapes.cuh -- analytic ray/primitive intersection for the nine Sail shapes, split B200-first
// into (1) a nearest-hit *scan* that computes only the hit distance of every object (what
// `Intersect.d` would be) and (2) a *hit record* evaluated once, for the winning object only.
//
// The reference evaluates the whole hit record -- uv, dpdu/dpdv, normal, surface colour, frame
// transforms -- for every object a ray touches and throws all but the nearest away
// (/root/reference/src/shader/shape/shader.shape.js:28-51).  Because the winner is chosen on `d` alone
// (strict `<`, ascending index, shader.shape.js:36) the two-phase form returns identical values.
// Shadow rays (src/shader/light/shader.light.js:24-31) only ever read `d`, so they run the scan only.
//
// Divisions by ray.dir share one reciprocal per ray (GLSL `/` is a*rcp(b), DESIGN.md "Arithmetic").
#ifndef SAIL_SHAPES_CUH
#define SAIL_SHAPES_CUH
#include "sail_scene.h"
#include "sail_vec.cuh"

#define SAIL_MAX_DISTANCE 1e5f
#define SAIL_EPSILON 1e-5f
#define SAIL_ONEMINUSEPSILON 0.9999f
#define SAIL_INF 1e5f
#define SAIL_PI 3.141592653589793f
#define SAIL_TWO_PI (2.0f * SAIL_PI)
#define SAIL_INVPI 0.3183098861837907f
#define SAIL_PIOVER2 1.570796326794896f
#define SAIL_PIOVER4 0.785398163397448f
// phi / (2.0*PI) and theta / PI multiply by the correctly rounded reciprocal of the FP32 constant.
#define SAIL_RCP_TWO_PI (1.0f / SAIL_TWO_PI)
#define SAIL_RCP_PI (1.0f / SAIL_PI)

struct Ray {
    V3 o, d;
};

// The scene as the kernels see it: the used part of the SailSceneBlock, packed back to back in dynamic
// shared memory -- [16 B header][n shapes][n surfaces][ln+1 lights] -- so a CTA reserves only what the
// scene needs (1.1 KB for the Cornell box) and 8 CTAs fit an SM next to the wavefront's path slots.
struct SceneView {
    const SailShape* shape;
    const SailSurface* surf;
    const SailLight* light;
    int n, ln;
    unsigned emissive;  // SailSceneBlock::emissiveMask
};
SD SceneView sceneViewAt(const unsigned char* base) {
    SceneView v;
    const int* hdr = reinterpret_cast<const int*>(base);
    v.n = hdr[0];
    v.ln = hdr[1];
    v.emissive = (unsigned)hdr[3];
    v.shape = reinterpret_cast<const SailShape*>(base + 16);
    v.surf = reinterpret_cast<const SailSurface*>(base + 16 + (size_t)v.n * sizeof(SailShape));
    v.light = reinterpret_cast<const SailLight*>(base + 16 + (size_t)v.n * (sizeof(SailShape) + sizeof(SailSurface)));
    return v;
}

// Kernel specialisation: the reference compiles only the plugins a scene uses into its trace program
// (/root/reference/src/scene/scene.js:70-112, src/shader/generator.js:107-123).  The AOT equivalent is one
// kernel instantiation per feature set F; code for absent features is not emitted (smaller I-cache
// footprint, fewer registers).  A variant is chosen from the *content* of the uploaded scene.
enum { SAIL_F_QUADRICS = 1,  // cone, cylinder, disk, hyperboloid, paraboloid present
       SAIL_F_GLOSSY = 2,    // metal or glass present
       SAIL_F_TEXTURED = 4,  // a uv-dependent texture (checkerboard, checkerboard2, bilerp, uvf) present
       SAIL_F_NEE = 8,       // ln > 0: next-event estimation + shadow rays
       SAIL_F_ALL = 15 };

// A scene-specialised kernel (sail_jit.cpp) also knows which shape / material / texture / light plugins the scene
// contains and compiles only those arms -- the reference's per-scene shader generation (generator.js:107-123).  In the
// ahead-of-time variants every arm the feature set F allows is present.
#ifdef SAIL_JIT_NRUNS
#define SAIL_HAS_SHAPE(c) ((((unsigned)(SAIL_JIT_SHAPES)) >> (c)) & 1u)
#define SAIL_HAS_MAT(c) ((((unsigned)(SAIL_JIT_MATS)) >> (c)) & 1u)
#define SAIL_HAS_TEX(c) ((((unsigned)(SAIL_JIT_TEXS)) >> (c)) & 1u)
#define SAIL_HAS_LIGHT(c) ((((unsigned)(SAIL_JIT_LIGHTS)) >> (c)) & 1u)
// shapes that are the geometry of an area light: only their samplers (sampleX / the stub + normalForX) are compiled
#define SAIL_HAS_LIGHT_SHAPE(c) ((((unsigned)(SAIL_JIT_LIGHT_SHAPES)) >> (c)) & 1u)
#define SAIL_HAS_STUB_LIGHT_SHAPE ((((unsigned)(SAIL_JIT_LIGHT_SHAPES)) & ~((1u << SAIL_SPHERE) | (1u << SAIL_RECTANGLE) | (1u << SAIL_DISK))) != 0u)
#define SAIL_HAS_SMOOTH_GLASS ((SAIL_JIT_GLASS) & 1)
#define SAIL_HAS_ROUGH_GLASS ((SAIL_JIT_GLASS) & 2)
#define SAIL_HAS_LAMBERT ((SAIL_JIT_MATTE) & 1)
#define SAIL_HAS_OREN_NAYAR ((SAIL_JIT_MATTE) & 2)
#define SAIL_HAS_TR_ISO ((SAIL_JIT_TR) & 1)
#define SAIL_HAS_TR_ANISO ((SAIL_JIT_TR) & 2)
#else
#define SAIL_HAS_LAMBERT 1
#define SAIL_HAS_OREN_NAYAR 1
#define SAIL_HAS_TR_ISO 1
#define SAIL_HAS_TR_ANISO 1
#define SAIL_HAS_SHAPE(c) 1
#define SAIL_HAS_MAT(c) 1
#define SAIL_HAS_TEX(c) 1
#define SAIL_HAS_LIGHT(c) 1
#define SAIL_HAS_LIGHT_SHAPE(c) 1
#define SAIL_HAS_STUB_LIGHT_SHAPE 1
#define SAIL_HAS_SMOOTH_GLASS 1
#define SAIL_HAS_ROUGH_GLASS 1
#endif

// Out-of-line transcendentals: one copy per kernel instead of one per call site (I-cache footprint).
static __device__ __noinline__ float sail_sin(float x) { return dm_sin(x); }
static __device__ __noinline__ float2 sail_sincos(float x) { float s, c; dm_sincos(x, &s, &c); return make_float2(s, c); }
static __device__ __noinline__ float sail_acos(float x) { return dm_acos(x); }
static __device__ __noinline__ float sail_atan2(float y, float x) { return dm_atan2(y, x); }
// Single-use sites of the small variants are inlined instead (no call overhead, code still fits the I-cache).
#ifdef SAIL_INLINE_TRANSCENDENTALS  // sail_jit.cpp / SAIL_B200_JIT_DEFINES: inline them in a big kernel too
#define SAIL_SMALL(F) 1
#else
#define SAIL_SMALL(F) (!((F) & (SAIL_F_GLOSSY | SAIL_F_QUADRICS | SAIL_F_NEE)))
#endif
// dm_sincos for |x| < 2^22 (the caller passes 2*pi*u, u in [0, 1]): its out-of-range arm is unreachable and dropped, and the
// sine and cosine polynomials, which share z = r*r, advance as one packed pair (FFMA2 / FMUL2).  Every operation and its
// order are those of dm_sin_kernel / dm_cos_kernel, so the result equals dm_sincos(x) bit for bit
// (tests/test_gpu_parity.py::test_unit_domain_math).
static __constant__ float2 sailSinCosC0 = {-1.9515295891e-4f, 2.443315711809948e-5f};
static __constant__ float2 sailSinCosC1 = {8.3321608736e-3f, -1.388731625493765e-3f};
static __constant__ float2 sailSinCosC2 = {-1.6666654611e-1f, 4.166664568298827e-2f};
SD float2 sincosUnit(float x) {
    const float n = rintf(dm_mul(x, DM_TWO_OVER_PI));
    float r = dm_fma(-n, DM_PIO2_HI, x);
    r = dm_fma(-n, DM_PIO2_MID, r);
    r = dm_fma(-n, DM_PIO2_LO, r);
    const int q = (int)n;
    const float z = dm_mul(r, r);
    const F2 zz = splat2(z);
    F2 p = fma2(*reinterpret_cast<const F2*>(&sailSinCosC0), zz, *reinterpret_cast<const F2*>(&sailSinCosC1));
    p = fma2(p, zz, *reinterpret_cast<const F2*>(&sailSinCosC2));
    p = mul2(p, zz);
    float psz, pcz;
    upk2(p, psz, pcz);
    const float s = dm_fma(psz, r, r);
    const float c = dm_add(dm_fma(-0.5f, z, 1.0f), dm_mul(pcz, z));
    const float vs = (q & 1) ? c : s, vc = (q & 1) ? s : c;
    return make_float2((q & 2) ? -vs : vs, ((q + 1) & 2) ? -vc : vc);
}
template <int F> SD float2 sincosF(float x) {
    if (SAIL_SMALL(F)) return sincosUnit(x);
    return sail_sincos(x);
}
// theta = acos(..) lies in [0, pi] (or is NaN): inside sincosUnit's domain; the unused cosine select is dropped by the compiler
template <int F> SD float sinF(float x) { return SAIL_SMALL(F) ? sincosUnit(x).x : sail_sin(x); }
template <int F> SD float acosF(float x) { return SAIL_SMALL(F) ? dm_acos(x) : sail_acos(x); }

// The winning hit (const/struct.glsl:1-13 minus the dead dpdv / seed fields).
struct Hit {
    float d;
    int index;       // object id, -1 on a miss (the HIT_ID AOV)
    V3 p, n, dpdu;   // world space; n already flipped toward the ray (shader.shape.js:39-40)
    V3 sc, emission;
    int matCategory;
    bool into;
    bool unitDpdu;   // dpdu is an exact axis unit vector (box faces): normalize() is the identity
};

// ---- util/utility.glsl:37-51 -----------------------------------------------------------------
SD bool quadratic(float A, float B, float C, float& t0, float& t1) {
    float discrim = B * B - 4.0f * A * C;
    if (discrim < 0.0f) return false;
    float rootDiscrim = dm_sqrt(discrim);
    float q;
    if (B < 0.0f) q = -0.5f * (B - rootDiscrim);
    else q = -0.5f * (B + rootDiscrim);
    t0 = fdiv(q, A);
    t1 = fdiv(C, q);
    if (t0 > t1) { float tmp = t0; t0 = t1; t1 = tmp; }
    return true;
}

// ---- shape/boundbox.glsl:6-17 and the slab core of cube.glsl:65-74 / cornellbox.glsl:67-76 ----
// tMin = (box.min - o) / d, tMax = (box.max - o) / d per axis, then the component-wise min / max.  Which corner the
// shader calls min or max does not matter (min/max are symmetric), so each axis is evaluated as ONE packed pair
// {lo, hi}: one FADD2 (subtract the origin) and one FMUL2 (times 1/d) instead of two FADDs and two FMULs.
SD void slab(const SailShape& s, V3 o, V3 rd, float& tNear, float& tFar) {
    F2 tx = mul2(sub2(ld2(s.lohi), pk2(o.x, o.x)), pk2(rd.x, rd.x));
    F2 ty = mul2(sub2(ld2(s.lohi + 2), pk2(o.y, o.y)), pk2(rd.y, rd.y));
    F2 tz = mul2(sub2(ld2(s.lohi + 4), pk2(o.z, o.z)), pk2(rd.z, rd.z));
    float ax, bx, ay, by, az, bz;
    upk2(tx, ax, bx); upk2(ty, ay, by); upk2(tz, az, bz);
    tNear = fmax2(fmax2(fmin2(ax, bx), fmin2(ay, by)), fmin2(az, bz));
    tFar = fmin2(fmin2(fmax2(ax, bx), fmax2(ay, by)), fmax2(az, bz));
}
SD V3 loOf(const SailShape& s) { return mk3(s.lohi[0], s.lohi[2], s.lohi[4]); }
SD V3 hiOf(const SailShape& s) { return mk3(s.lohi[1], s.lohi[3], s.lohi[5]); }
// testBoundbox with Boundbox(lo, hi) (the constructor order (max, min) is immaterial, see above).
SD bool boundsHit(const SailShape& s, V3 o, V3 rd) {
    float tNear, tFar;
    slab(s, o, rd, tNear, tFar);
    if (tNear < 0.0f && tFar < 0.0f) return false;
    return tNear < tFar;
}

// Quadric tail shared by cone / cylinder / hyperboloid / paraboloid (e.g. cone.glsl:62-79):
// choose the root, retry with the far root when the near one is outside [zLo, zHi].
SD float quadricRoot(float a, float b, float c, V3 ol, V3 dl, float zLo, float zHi) {
    float t1, t2;
    if (!quadratic(a, b, c, t1, t2)) return SAIL_MAX_DISTANCE;
    if (t2 < -SAIL_EPSILON) return SAIL_MAX_DISTANCE;
    float t = t1;
    if (t1 < SAIL_EPSILON) t = t2;
    float hz = ol.z + t * dl.z;
    if (hz < zLo || hz > zHi) {
        if (t == t2) return SAIL_MAX_DISTANCE;
        t = t2;
        hz = ol.z + t * dl.z;
        if (hz < zLo || hz > zHi) return SAIL_MAX_DISTANCE;
    }
    if (t >= SAIL_MAX_DISTANCE) return SAIL_MAX_DISTANCE;
    return t;
}

SD int opaqueInt(int v) {
    asm("" : "+r"(v));
    return v;
}
// Shapes whose intersectX() starts with testBoundboxForX() (sphere.glsl:10-16, cone / cylinder / hyperboloid / paraboloid alike)
#define SAIL_BOUNDED_SHAPE(c) ((c) == SAIL_SPHERE || (c) == SAIL_CONE || (c) == SAIL_CYLINDER || (c) == SAIL_HYPERBOLOID || (c) == SAIL_PARABOLOID)
// Runs shorter than this are unrolled object by object (measured: two spheres are faster unrolled, config 5 -1.3 % otherwise).
#define SAIL_JIT_UNROLL_BELOW(c) 3
// Phase 1: the value intersectX() would leave in `result.d` (MAX_DISTANCE = no hit).
template <int F>
SD float shapeDistance(const SailShape& s, const Ray& ray, V3 rd, const int fixedCat = -1, const bool boundsDone = false) {
    const float* q = s.q;
    // The three common shapes are tested by a compare chain (2-3 instructions each); a `switch` over all
    // categories compiles to a bounds check + jump-table load + indirect branch, 11 instructions per object.
    // opaqueInt() keeps the compiler from folding the chain back into that switch.
    // fixedCat >= 0: the category is a compile-time constant of a scene-specialised kernel (SAIL_JIT_NRUNS below)
    const int cat = fixedCat >= 0 ? fixedCat : s.category;
    if (cat == SAIL_CUBE || cat == SAIL_CORNELLBOX) {
        // cube.glsl:65-79: the near slab hit if it is in front, else the far one
        // cornellbox.glsl:67-81: always the far slab hit
        float tNear, tFar;
        slab(s, ray.o, rd, tNear, tFar);
        float t = -1.0f;
        if (tNear < tFar) t = (cat == SAIL_CUBE && tNear > SAIL_EPSILON) ? tNear : tFar;
        return (t > SAIL_EPSILON) ? t : SAIL_MAX_DISTANCE;
    }
    const int cat2 = fixedCat >= 0 ? cat : opaqueInt(cat);
    switch (cat2 == SAIL_SPHERE ? SAIL_SPHERE : (cat2 == SAIL_RECTANGLE ? SAIL_RECTANGLE : ((F & SAIL_F_QUADRICS) ? cat : 0))) {
        case SAIL_SPHERE: if (SAIL_HAS_SHAPE(SAIL_SPHERE)) {  // sphere.glsl:10-16,45-62
            if (!boundsDone && !boundsHit(s, ray.o, rd)) return SAIL_MAX_DISTANCE;
            // dl = objToLocal(d) = (-d.z, d.x, d.y), ol likewise from w = o - c.  A product of two negated factors equals the
            // product of the factors, so the nine products of dot(dl,dl), dot(ol,dl), dot(ol,ol) are formed from the world-space
            // components as they lie in registers (three packed multiplies, no swizzled copies) and summed in the local
            // order x, y, z = world z, x, y -- the same values in the same order as dot3() on the swizzled vectors.
            const V3 w = ray.o - ld3(q + SQ_SPH_C);
            const V3 dd = ray.d * ray.d, wd = w * ray.d, ww = w * w;
            float a = dd.z + dd.x + dd.y;
            float b = 2.0f * (wd.z + wd.x + wd.y);
            float c = ww.z + ww.x + ww.y - q[SQ_SPH_RR];
            float t1, t2;
            if (!quadratic(a, b, c, t1, t2)) return SAIL_MAX_DISTANCE;
            if (t2 < SAIL_EPSILON) return SAIL_MAX_DISTANCE;
            float t = t1;
            if (t1 < SAIL_EPSILON) t = t2;
            if (t >= SAIL_MAX_DISTANCE) return SAIL_MAX_DISTANCE;
            return t;
        }
        case SAIL_RECTANGLE: if (SAIL_HAS_SHAPE(SAIL_RECTANGLE)) {  // rectangle.glsl:32-55
            V3 nn = ld3(q + SQ_RECT_N), ss = ld3(q + SQ_RECT_SS), ts = ld3(q + SQ_RECT_TS);
            V3 dl = worldToLocal(ray.d, nn, ss, ts);
            V3 ol = worldToLocal(ray.o - loOf(s), nn, ss, ts);
            if (dl.z == 0.0f) return SAIL_MAX_DISTANCE;
            float t = fdiv(-ol.z, dl.z);
            if (t < SAIL_EPSILON) return SAIL_MAX_DISTANCE;
            float hx = ol.x + t * dl.x, hy = ol.y + t * dl.y;
            if (hx > q[SQ_RECT_MAXX] || hy > q[SQ_RECT_MAXY] || hx < -SAIL_EPSILON || hy < -SAIL_EPSILON)
                return SAIL_MAX_DISTANCE;
            return t;
        }
        case SAIL_CONE: if ((F & SAIL_F_QUADRICS) && SAIL_HAS_SHAPE(SAIL_CONE)) {  // cone.glsl:11-17,48-81
            if (!boundsDone && !boundsHit(s, ray.o, rd)) return SAIL_MAX_DISTANCE;
            V3 dl = objToLocal(ray.d), ol = objToLocal(ray.o - ld3(q + SQ_CONE_P));
            float k = q[SQ_CONE_K], h = q[SQ_CONE_H];
            float a = dl.x * dl.x + dl.y * dl.y - k * dl.z * dl.z;
            float b = 2.0f * (dl.x * ol.x + dl.y * ol.y - k * dl.z * (ol.z - h));
            float c = ol.x * ol.x + ol.y * ol.y - k * (ol.z - h) * (ol.z - h);
            return quadricRoot(a, b, c, ol, dl, -SAIL_EPSILON, h);
        }
        case SAIL_CYLINDER: if ((F & SAIL_F_QUADRICS) && SAIL_HAS_SHAPE(SAIL_CYLINDER)) {  // cylinder.glsl:11-17,40-72
            if (!boundsDone && !boundsHit(s, ray.o, rd)) return SAIL_MAX_DISTANCE;
            V3 dl = objToLocal(ray.d), ol = objToLocal(ray.o - ld3(q + SQ_CYL_P));
            float a = dl.x * dl.x + dl.y * dl.y;
            float b = 2.0f * (dl.x * ol.x + dl.y * ol.y);
            float c = ol.x * ol.x + ol.y * ol.y - q[SQ_CYL_RR];
            return quadricRoot(a, b, c, ol, dl, -SAIL_EPSILON, q[SQ_CYL_H]);
        }
        case SAIL_DISK: if ((F & SAIL_F_QUADRICS) && SAIL_HAS_SHAPE(SAIL_DISK)) {  // disk.glsl:36-53
            V3 dl = objToLocal(ray.d), ol = objToLocal(ray.o - ld3(q + SQ_DISK_P));
            if (dl.z == 0.0f) return SAIL_MAX_DISTANCE;
            float t = fdiv(-ol.z, dl.z);
            if (t <= 0.0f) return SAIL_MAX_DISTANCE;
            float hx = ol.x + t * dl.x, hy = ol.y + t * dl.y;
            float dist2 = hx * hx + hy * hy;
            if (dist2 > q[SQ_DISK_RR] || dist2 < q[SQ_DISK_IRR]) return SAIL_MAX_DISTANCE;
            if (t >= SAIL_MAX_DISTANCE) return SAIL_MAX_DISTANCE;
            return t;
        }
        case SAIL_HYPERBOLOID: if ((F & SAIL_F_QUADRICS) && SAIL_HAS_SHAPE(SAIL_HYPERBOLOID)) {  // hyperboloid.glsl:13-24,60-90
            if (!boundsDone && !boundsHit(s, ray.o, rd)) return SAIL_MAX_DISTANCE;
            V3 dl = objToLocal(ray.d), ol = objToLocal(ray.o - ld3(q + SQ_HYP_P));
            float ah = q[SQ_HYP_AH], ch = q[SQ_HYP_CH];
            float a = ah * dl.x * dl.x + ah * dl.y * dl.y - ch * dl.z * dl.z;
            float b = 2.0f * (ah * dl.x * ol.x + ah * dl.y * ol.y - ch * dl.z * ol.z);
            float c = ah * ol.x * ol.x + ah * ol.y * ol.y - ch * ol.z * ol.z - 1.0f;
            return quadricRoot(a, b, c, ol, dl, q[SQ_HYP_ZMIN], q[SQ_HYP_ZMAX]);
        }
        case SAIL_PARABOLOID: if ((F & SAIL_F_QUADRICS) && SAIL_HAS_SHAPE(SAIL_PARABOLOID)) {  // paraboloid.glsl:12-20,51-82
            if (!boundsDone && !boundsHit(s, ray.o, rd)) return SAIL_MAX_DISTANCE;
            V3 dl = objToLocal(ray.d), ol = objToLocal(ray.o - ld3(q + SQ_PAR_P));
            float k = q[SQ_PAR_K];
            float a = k * (dl.x * dl.x + dl.y * dl.y);
            float b = 2.0f * k * (dl.x * ol.x + dl.y * ol.y) - dl.z;
            float c = k * (ol.x * ol.x + ol.y * ol.y) - ol.z;
            return quadricRoot(a, b, c, ol, dl, q[SQ_PAR_ZMIN], q[SQ_PAR_ZMAX]);
        }
    }
    return SAIL_MAX_DISTANCE;
}

// dm_rcp (detmath.h) of the three components of a direction: the two Newton FMAs of x and y issue as one packed
// pair; every lane computes exactly what dm_rcp computes.
SD V3 rcp3(V3 d) {
    float yx, yy;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(yx) : "f"(d.x));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(yy) : "f"(d.y));
    const F2 y0 = pk2(yx, yy);
    const F2 e = fma2(pk2(-d.x, -d.y), y0, splat2(1.0f));
    F2 y1;
    asm("fma.rn.ftz.f32x2 %0, %1, %2, %1;" : "=l"(y1) : "l"(y0), "l"(e));
    float rx, ry;
    upk2(y1, rx, ry);
    return mk3((rx != rx) ? yx : rx, (ry != ry) ? yy : ry, dm_rcp(d.z));
}

// shader.shape.js:28-37: nearest object by strict `<` in ascending index.  Returns (ins.d, object index
// or -1).  Out of line (one copy serves primary/bounce rays and shadow rays); everything travels in
// registers -- the scene block is addressed by its shared-memory window offset.
struct Nearest {
    float t;
    int id;
};
template <int F>
SD Nearest nearestObjectInline(const SceneView& sc, const Ray& ray) {
    V3 rd = rcp3(ray.d);
    Nearest r;
    r.t = SAIL_MAX_DISTANCE;
    r.id = -1;
#ifdef SAIL_JIT_NRUNS
    // Scene-specialised kernel (sail_jit.cpp): the object list is known at compile time as runs of equal shape
    // category -- what the reference gets from generating its shader per scene (generator.js:107-123).  Runs are
    // unrolled (no dispatch, no loop control, tests of different objects interleave); a run of three or more equal
    // shapes stays a loop over that one shape's code, so the kernel does not grow with the object count.
    constexpr int kRunCat[SAIL_JIT_NRUNS] = {SAIL_JIT_RUN_CATS};
    constexpr int kRunStart[SAIL_JIT_NRUNS] = {SAIL_JIT_RUN_STARTS};
    constexpr int kRunLen[SAIL_JIT_NRUNS] = {SAIL_JIT_RUN_LENS};
#pragma unroll
    for (int k = 0; k < SAIL_JIT_NRUNS; k++) {
        if (kRunLen[k] < SAIL_JIT_UNROLL_BELOW(kRunCat[k])) {
#pragma unroll
            for (int j = 0; j < kRunLen[k]; j++) {
                float t = shapeDistance<F>(sc.shape[kRunStart[k] + j], ray, rd, kRunCat[k]);
                if (t < r.t) { r.t = t; r.id = kRunStart[k] + j; }
            }
        } else if (SAIL_BOUNDED_SHAPE(kRunCat[k]) && kRunLen[k] <= 32) {
            // A run of equal shapes that test their bounding box first (sphere.glsl:10-16 ...): all the (cheap, uniform)
            // box tests first, then every lane walks ITS OWN list of boxes it passed, in ascending index.  One pass through
            // the quadratic serves different objects on different lanes, so the warp makes max-over-lanes passes (typically
            // 1-2) instead of one per object with a few lanes each.  Per object the same two steps in the same order, and the
            // running minimum still sees the objects in ascending index: identical result.
            unsigned pend = 0;
#pragma unroll 1
            for (int j = 0; j < kRunLen[k]; j++)
                if (boundsHit(sc.shape[kRunStart[k] + j], ray.o, rd)) pend |= 1u << j;
            while (pend) {
                const int i = kRunStart[k] + (__ffs(pend) - 1);
                pend &= pend - 1u;
                float t = shapeDistance<F>(sc.shape[i], ray, rd, kRunCat[k], true);
                if (t < r.t) { r.t = t; r.id = i; }
            }
        } else {
#pragma unroll 1
            for (int i = kRunStart[k]; i < kRunStart[k] + kRunLen[k]; i++) {
                float t = shapeDistance<F>(sc.shape[i], ray, rd, kRunCat[k]);
                if (t < r.t) { r.t = t; r.id = i; }
            }
        }
    }
#else
    const int n = sc.n;
    for (int i = 0; i < n; i++) {
        float t = shapeDistance<F>(sc.shape[i], ray, rd);
        if (t < r.t) { r.t = t; r.id = i; }
    }
#endif
    return r;
}
// Does the ray reach the surface of ANY object selected by `mask` (bit = object index; objects >= 32 always count)?
// Used on the last bounce of kernels without NEE, where only a hit on an emissive object can change the sample
// (sail_shade.cuh tracePath): when no selected object is touched at all, the nearest hit -- whatever it is -- adds
// nothing, and the scan of the other objects is skipped.  Same distance code as the scan, so no hit is ever missed.
template <int F>
SD bool touchesMasked(const SceneView& sc, const Ray& ray, unsigned mask) {
    V3 rd = rcp3(ray.d);
    bool any = false;
#ifdef SAIL_JIT_NRUNS
    constexpr int kRunCat[SAIL_JIT_NRUNS] = {SAIL_JIT_RUN_CATS};
    constexpr int kRunStart[SAIL_JIT_NRUNS] = {SAIL_JIT_RUN_STARTS};
    constexpr int kRunLen[SAIL_JIT_NRUNS] = {SAIL_JIT_RUN_LENS};
#pragma unroll
    for (int k = 0; k < SAIL_JIT_NRUNS; k++) {
#pragma unroll 1
        for (int i = kRunStart[k]; i < kRunStart[k] + kRunLen[k]; i++) {
            if (i < 32 && !((mask >> i) & 1u)) continue;
            any |= shapeDistance<F>(sc.shape[i], ray, rd, kRunCat[k]) < SAIL_MAX_DISTANCE;
        }
    }
#else
    const int n = sc.n;
    for (int i = 0; i < n; i++) {
        if (i < 32 && !((mask >> i) & 1u)) continue;
        any |= shapeDistance<F>(sc.shape[i], ray, rd) < SAIL_MAX_DISTANCE;
    }
#endif
    return any;
}
template <int F>
static __device__ __noinline__ Nearest nearestObjectRegs(unsigned scAddr, float ox, float oy, float oz, float dx, float dy, float dz) {
    const SceneView sc = sceneViewAt(reinterpret_cast<const unsigned char*>(__cvta_shared_to_generic(scAddr)));
    Ray ray;
    ray.o = mk3(ox, oy, oz);
    ray.d = mk3(dx, dy, dz);
    return nearestObjectInline<F>(sc, ray);
}
// Variants without next-event estimation have a single call site (the bounce ray): inline it.  With NEE the
// shadow rays are a second site and one out-of-line copy serves both (I-cache footprint).
template <int F>
SD Nearest nearestObject(const SceneView& sc, const Ray& ray) {
#ifdef SAIL_INLINE_SCAN  // A/B knob: two inlined copies (path ray, shadow ray) instead of one shared out-of-line copy
    return nearestObjectInline<F>(sc, ray);
#endif
    if (!(F & SAIL_F_NEE)) return nearestObjectInline<F>(sc, ray);
    return nearestObjectRegs<F>((unsigned)__cvta_generic_to_shared(reinterpret_cast<const unsigned char*>(sc.shape) - 16), ray.o.x, ray.o.y, ray.o.z, ray.d.x, ray.d.y, ray.d.z);
}

// ---- procedural textures (texture/shader.texture.js:22-29) -------------------------------------
SD bool textureNeedsUV(int texCategory) {
    return texCategory == SAIL_CHECKERBOARD || texCategory == SAIL_CHECKERBOARD2 || texCategory == SAIL_BILERP ||
           texCategory == SAIL_UVF;
}
template <int F>
SD V3 surfaceColor(const SailSurface& u, V2 uv) {
    const float* t = u.t;
    if (F & SAIL_F_TEXTURED) switch (u.texCategory) {
        case SAIL_CHECKERBOARD: if (SAIL_HAS_TEX(SAIL_CHECKERBOARD)) {  // checkerboard.glsl:6-20
            float ax = uv.x * t[13], ay = uv.y * t[13], width = t[14];
            float fx = ax - dm_floor(ax), fy = ay - dm_floor(ay);
            bool in_outline = (fx < width || fx > 1.0f - width) || (fy < width || fy > 1.0f - width);
            return in_outline ? splat3(0.5f) : splat3(1.0f);
        }
        case SAIL_CHECKERBOARD2: if (SAIL_HAS_TEX(SAIL_CHECKERBOARD2)) {  // checkerboard2.glsl:7-16
            float ux = dm_floor(uv.x * t[13]), uy = dm_floor(uv.y * t[13]);
            return (toInt(ux + uy) % 2 == 0) ? ld3(t + ST_C1) : ld3(t + ST_C2);
        }
        case SAIL_BILERP: if (SAIL_HAS_TEX(SAIL_BILERP)) {  // bilerp.glsl:1-14 (broken in the reference, SURVEY Q12: evident intent)
            float ux = uv.x, uy = uv.y;
            return (1.0f - ux) * (1.0f - uy) * ld3(t + ST_B00) + (1.0f - ux) * uy * ld3(t + ST_B01) +
                   ux * (1.0f - uy) * ld3(t + ST_B10) + ux * uy * ld3(t + ST_B11);
        }
        case SAIL_UVF:  // uvf.glsl:1-3
            return mk3(uv.x - dm_floor(uv.x), uv.y - dm_floor(uv.y), 0.0f);
    }
    return ld3(t + ST_RGB);  // UNIFORM_COLOR, and MIXF / SCALE folded to a constant at upload
}

SD float azimuthU(float hx, float hy) {  // phi = atan(y,x); if (phi<0) phi += 2PI; u = phi/(2PI)
    float phi = sail_atan2(hy, hx);
    if (phi < 0.0f) phi += SAIL_TWO_PI;
    return phi * SAIL_RCP_TWO_PI;
}
SD V3 dpduAround(float hx, float hy) { return mk3(-2.0f * SAIL_PI * hy, 2.0f * SAIL_PI * hx, 0.0f); }

// cube.glsl:25-38 / cornellbox.glsl:39-51 face selection
SD V3 boxFaceNormal(V3 hit, V3 mn, V3 mx, float c) {
    if (hit.x < mn.x + 0.0001f) return mk3(c * -1.0f, c * 0.0f, c * 0.0f);
    else if (hit.x > mx.x - 0.0001f) return mk3(c * 1.0f, c * 0.0f, c * 0.0f);
    else if (hit.y < mn.y + 0.0001f) return mk3(c * 0.0f, c * -1.0f, c * 0.0f);
    else if (hit.y > mx.y - 0.0001f) return mk3(c * 0.0f, c * 1.0f, c * 0.0f);
    else if (hit.z < mn.z + 0.0001f) return mk3(c * 0.0f, c * 0.0f, c * -1.0f);
    else return mk3(c * 0.0f, c * 0.0f, c * 1.0f);
}
SD V3 boxDpdu(V3 normal) {  // cube.glsl:40-48
    if (dm_abs(normal.x) < 0.5f) return cross3(normal, mk3(1.0f, 0.0f, 0.0f));
    return cross3(normal, mk3(0.0f, 1.0f, 0.0f));
}

// Phase 2: the full hit record of object `i` at distance `t` (the body of intersectX() after the hit
// test, the per-arm emission mask of shader.shape.js:47-49 and the tail of :38-40).
template <int F>
SD void hitRecord(const SceneView& sc, int i, float t, const Ray& ray, Hit& h) {
    const SailShape& s = sc.shape[i];
    const SailSurface& u = sc.surf[i];
    const float* q = s.q;
    const bool wantUV = (F & SAIL_F_TEXTURED) && textureNeedsUV(u.texCategory);
    V3 gn = splat3(0.0f);  // geometric normal as intersectX() returns it
    V2 uv = mk2(0.0f, 0.0f);
    h.d = t;
    h.index = i;
    h.unitDpdu = s.category == SAIL_CUBE || s.category == SAIL_CORNELLBOX;
    const int cat = s.category;
    switch (cat) {
        case SAIL_CUBE:          // cube.glsl:25-48,79-85
        case SAIL_CORNELLBOX: {  // cornellbox.glsl:23-51,81-88
            // One branch-free path for both boxes: the face is picked with the shader's comparison chain
            // (x-, x+, y-, y+, z-, z+), its normal and dpdu come from the per-face table built at upload.
            V3 mn = loOf(s), mx = hiOf(s);
            h.p = ray.o + t * ray.d;
            int face = 5;
            face = (h.p.z < mn.z + 0.0001f) ? 4 : face;
            face = (h.p.y > mx.y - 0.0001f) ? 3 : face;
            face = (h.p.y < mn.y + 0.0001f) ? 2 : face;
            face = (h.p.x > mx.x - 0.0001f) ? 1 : face;
            face = (h.p.x < mn.x + 0.0001f) ? 0 : face;
            const float* ft = q + SQ_BOX_FACE + face * SQ_BOX_FACE_STRIDE;
            gn = ld3(ft);
            h.dpdu = ld3(ft + 3);
            if (s.category == SAIL_CORNELLBOX) {  // getCornellboxColor, cornellbox.glsl:23-37
                V3 wall = (face == 0) ? mk3(0.25f, 0.75f, 0.25f) : mk3(0.25f, 0.25f, 0.75f);  // GREEN : BLUE
                float rest = (face <= 3 || h.p.z > mn.z + 0.0001f) ? 1.0f : 0.0f;               // WHITE : BLACK
                h.sc = (face <= 1) ? wall : splat3(rest);
            } else {
                if (wantUV) {  // getCubeUV, cube.glsl:54-63 (compares the *shifted* hit, as written)
                    V3 tr = mx - mn, hs = h.p - mn;
                    if (hs.x < mn.x + 0.0001f || hs.x > mx.x - 0.0001f) uv = mk2(fdiv(hs.y, tr.y), fdiv(hs.z, tr.z));
                    else if (hs.y < mn.y + 0.0001f || hs.y > mx.y - 0.0001f) uv = mk2(fdiv(hs.x, tr.x), fdiv(hs.z, tr.z));
                    else uv = mk2(fdiv(hs.x, tr.x), fdiv(hs.y, tr.y));
                }
                h.sc = surfaceColor<F>(u, uv);
            }
            break;
        }
        case SAIL_SPHERE: if (SAIL_HAS_SHAPE(SAIL_SPHERE)) {  // sphere.glsl:33-43,62-84
            V3 c = ld3(q + SQ_SPH_C);
            // ol + t*dl with ol, dl the swizzles (-z, x, y) of w = o - c and d: negation commutes with every rounding, so this is
            // the swizzle of w + t*d
            V3 hl = objToLocal((ray.o - c) + t * ray.d);
            float r = q[SQ_SPH_R];
            float theta = acosF<F>(fclamp(hl.z * q[SQ_SPH_INVR], -1.0f, 1.0f));
            if (wantUV) {
                float gx = hl.x;
                if (hl.x == 0.0f && hl.y == 0.0f) gx = 1e-5f * r;  // guards phi only (Q18)
                uv = mk2(azimuthU(gx, hl.y), theta * SAIL_RCP_PI);
            }
            float zRadius = dm_sqrt(hl.x * hl.x + hl.y * hl.y);
            float invZRadius = dm_rcp(zRadius);
            float cosPhi = hl.x * invZRadius, sinPhi = hl.y * invZRadius;
            V3 dpdu = dpduAround(hl.x, hl.y);
            V3 dpdv = SAIL_PI * mk3(hl.z * cosPhi, hl.z * sinPhi, -r * sinF<F>(theta));
            V3 nl = normalize3(cross3(dpdv, dpdu));
            h.sc = surfaceColor<F>(u, uv);
            h.p = objToWorld(hl) + c;
            gn = objToWorld(nl);
            h.dpdu = objToWorld(dpdu);
            break;
        }
        case SAIL_RECTANGLE: if (SAIL_HAS_SHAPE(SAIL_RECTANGLE)) {  // rectangle.glsl:36-62
            V3 nn = ld3(q + SQ_RECT_N), ss = ld3(q + SQ_RECT_SS), ts = ld3(q + SQ_RECT_TS), mn = loOf(s);
            V3 dl = worldToLocal(ray.d, nn, ss, ts);
            V3 ol = worldToLocal(ray.o - mn, nn, ss, ts);
            V3 hl = ol + t * dl;
            if (wantUV) uv = mk2(hl.x * q[SQ_RECT_INVMAXX], hl.y * q[SQ_RECT_INVMAXY]);
            h.sc = surfaceColor<F>(u, uv);
            h.p = localToWorld(hl, nn, ss, ts) + mn;
            gn = nn;
            h.dpdu = ld3(q + SQ_RECT_DPDU);
            break;
        }
        case SAIL_CONE: if ((F & SAIL_F_QUADRICS) && SAIL_HAS_SHAPE(SAIL_CONE)) {  // cone.glsl:31-36,83-98
            V3 p = ld3(q + SQ_CONE_P);
            V3 dl = objToLocal(ray.d), ol = objToLocal(ray.o - p);
            V3 hl = ol + t * dl;
            float v = hl.z * q[SQ_CONE_INVH];
            if (wantUV) uv = mk2(azimuthU(hl.x, hl.y), v);
            V3 dpdu = dpduAround(hl.x, hl.y);
            float inv = dm_rcp(1.0f - v);
            V3 dpdv = mk3(-hl.x * inv, -hl.y * inv, q[SQ_CONE_H]);
            V3 nl = normalize3(cross3(dpdu, dpdv));
            h.sc = surfaceColor<F>(u, uv);
            h.p = objToWorld(hl) + p;
            gn = objToWorld(nl);
            h.dpdu = objToWorld(dpdu);
            break;
        }
        case SAIL_CYLINDER: if ((F & SAIL_F_QUADRICS) && SAIL_HAS_SHAPE(SAIL_CYLINDER)) {  // cylinder.glsl:31-34,74-89
            V3 p = ld3(q + SQ_CYL_P);
            V3 dl = objToLocal(ray.d), ol = objToLocal(ray.o - p);
            V3 hl = ol + t * dl;
            if (wantUV) uv = mk2(azimuthU(hl.x, hl.y), hl.z * q[SQ_CYL_INVH]);
            V3 dpdu = dpduAround(hl.x, hl.y);
            V3 dpdv = mk3(0.0f, 0.0f, q[SQ_CYL_H]);
            V3 nl = normalize3(cross3(dpdu, dpdv));
            h.sc = surfaceColor<F>(u, uv);
            h.p = objToWorld(hl) + p;
            gn = objToWorld(nl);
            h.dpdu = objToWorld(dpdu);
            break;
        }
        case SAIL_DISK: if ((F & SAIL_F_QUADRICS) && SAIL_HAS_SHAPE(SAIL_DISK)) {  // disk.glsl:27-30,55-74
            V3 p = ld3(q + SQ_DISK_P);
            V3 dl = objToLocal(ray.d), ol = objToLocal(ray.o - p);
            V3 hl = ol + t * dl;
            float dist2 = hl.x * hl.x + hl.y * hl.y;
            float rHit = dm_sqrt(dist2);
            if (wantUV) {
                float oneMinusV = (rHit - q[SQ_DISK_INNER]) * q[SQ_DISK_INVDR];
                uv = mk2(azimuthU(hl.x, hl.y), 1.0f - oneMinusV);
            }
            V3 dpdu = dpduAround(hl.x, hl.y);
            V3 dpdv = vdivs(mk3(hl.x, hl.y, 0.0f) * q[SQ_DISK_DIFF], rHit);
            V3 nl = normalize3(cross3(dpdu, dpdv));
            h.sc = surfaceColor<F>(u, uv);
            h.p = objToWorld(hl) + p;
            gn = objToWorld(nl);
            h.dpdu = objToWorld(dpdu);
            break;
        }
        case SAIL_HYPERBOLOID: if ((F & SAIL_F_QUADRICS) && SAIL_HAS_SHAPE(SAIL_HYPERBOLOID)) {  // hyperboloid.glsl:41-46,92-110
            V3 p = ld3(q + SQ_HYP_P), p1 = ld3(q + SQ_HYP_P1), p2 = ld3(q + SQ_HYP_P2);
            V3 dl = objToLocal(ray.d), ol = objToLocal(ray.o - p);
            V3 hl = ol + t * dl;
            float v = (hl.z - p1.z) * q[SQ_HYP_INVDZ];
            V3 pr = (1.0f - v) * p1 + v * p2;
            float phi = sail_atan2(pr.x * hl.y - hl.x * pr.y, hl.x * pr.x + hl.y * pr.y);
            if (phi < 0.0f) phi += SAIL_TWO_PI;
            uv = mk2(phi * SAIL_RCP_TWO_PI, v);
            float2 scp = sail_sincos(phi);
            float sinPhi = scp.x, cosPhi = scp.y;
            V3 dpdu = dpduAround(hl.x, hl.y);
            V3 dpdv = mk3((p2.x - p1.x) * cosPhi - (p2.y - p1.y) * sinPhi, (p2.x - p1.x) * sinPhi + (p2.y - p1.y) * cosPhi,
                          p2.z - p1.z);
            V3 nl = normalize3(cross3(dpdu, dpdv));
            h.sc = surfaceColor<F>(u, uv);
            h.p = objToWorld(hl) + p;
            gn = objToWorld(nl);
            h.dpdu = objToWorld(dpdu);
            break;
        }
        default: if (F & SAIL_F_QUADRICS) {  // SAIL_PARABOLOID: paraboloid.glsl:35-40,84-102
            V3 p = ld3(q + SQ_PAR_P);
            V3 dl = objToLocal(ray.d), ol = objToLocal(ray.o - p);
            V3 hl = ol + t * dl;
            if (wantUV) uv = mk2(azimuthU(hl.x, hl.y), (hl.z - q[SQ_PAR_ZMIN]) * q[SQ_PAR_INVDZ]);
            V3 dpdu = dpduAround(hl.x, hl.y);
            float inv = dm_rcp(2.0f * hl.z);
            V3 dpdv = q[SQ_PAR_DZ] * mk3(hl.x * inv, hl.y * inv, 1.0f);
            V3 nl = normalize3(cross3(dpdu, dpdv));
            h.sc = surfaceColor<F>(u, uv);
            h.p = objToWorld(hl) + p;
            gn = objToWorld(nl);
            h.dpdu = objToWorld(dpdu);
            break;
        }
    }
    // shader.shape.js:47-49: emission only when the ray faces the (possibly reversed) normal.
    V3 fn = s.rev * gn;
    bool faceObj = dot3(fn, ray.d) < -SAIL_EPSILON;
    h.emission = faceObj ? ld3(s.emission) : splat3(0.0f);
    // shader.shape.js:38-40
    h.matCategory = s.matCategory;
    h.into = dot3(gn, ray.d) < -SAIL_EPSILON;
    h.n = h.into ? gn : -gn;
}

// intersectObjects() as one call (shader.shape.js:28-51).  Returns false on a miss.
template <int F>
SD bool intersectObjects(const SceneView& sc, const Ray& ray, Hit& h) {
    Nearest nr = nearestObject<F>(sc, ray);
    h.d = nr.t;
    h.index = nr.id;
    if (nr.id < 0) return false;
    hitRecord<F>(sc, nr.id, nr.t, ray, h);
    return true;
}

// testShadow(), shader.light.js:24-31: distance only.
template <int F>
SD bool testShadow(const SceneView& sc, V3 origin, V3 toLight) {
    Ray r;
    r.o = origin;
    r.d = toLight;
    float d = nearestObject<F>(sc, r).t;
    return d > SAIL_EPSILON && d < SAIL_ONEMINUSEPSILON;
}

#endif
