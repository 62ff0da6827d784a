// oracle/oracle.cpp -- TEST INFRASTRUCTURE ONLY: the CPU oracle for the Sail path-tracing hot path.
//
// A scalar FP32 restatement of the reference's GLSL (files cited in oracle/shader_*.h), consuming
// exactly the reference wire format: objects 18 x n, texParams 16 x tn, lights 18 x ln
// (src/core/tracer.js:42-90) plus the uniforms of src/core/shader.js:41-49.
//
// PARITY UNPINNED: the reference ships no tests, golden vectors or CPU implementation of this path
// and cannot be executed here (no JS engine, no graphics stack) -- SURVEY.md 4, 8(c).  The oracle's
// arithmetic model is stated in oracle/glsl.h.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load
// this library.  The product (libsail_b200.so) never links or calls it.
#include <omp.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <vector>

#ifdef SAIL_COUNT_OPS
thread_local uint64_t g_sail_ops = 0;
#endif

#include "shader_trace.h"
#include "oracle_filters.h"

static std::vector<float> g_objects, g_texparams, g_lights;
static int g_rr_start = 0;  // Russian roulette start bounce for the following frames (0 = off)

extern "C" {

void oracle_set_russian_roulette(int start_bounce) { g_rr_start = start_bounce > 0 ? start_bounce : 0; }

int oracle_set_scene(const float* objs, int n, const float* texparams, int tn, const float* lts, int ln) {
    if (n < 0 || tn < 0 || ln < 0) return -1;
    g_objects.assign(objs, objs + (size_t)n * 18);
    g_texparams.assign(texparams, texparams + (size_t)tn * 16);
    g_lights.assign(lts, lts + (size_t)ln * 18);
    U.n = n;
    U.tn = tn;
    U.ln = ln;
    U.objects_ = sampler2D{g_objects.data(), 18, n};
    U.texParams_ = sampler2D{g_texparams.data(), 16, tn};
    U.lights_ = sampler2D{g_lights.data(), 18, ln};
    return 0;
}

static void set_frame(const float* matrix, const float* eye, float textureWeight, float time, int maxBounces) {
    for (int i = 0; i < 16; i++) U.matrix.m[i] = real(matrix[i]);
    U.eye = vec3(eye[0], eye[1], eye[2]);
    U.textureWeight = real(textureWeight);
    U.timeSinceStart_ = real(time);
    U.maxBounces = maxBounces;
    U.rrStart = g_rr_start;
}

// One progressive frame (Tracer.render, src/core/tracer.js:92-101 + fstrace.glsl:6-17).
// accum: W*H*3 running mean, read as `cache` and overwritten with out_color.  Row 0 = bottom row.
// normal/position/hitId/depth may be NULL.  Returns the number of counted FP32 ops (0 unless built
// with SAIL_COUNT_OPS).
uint64_t oracle_render_frame(int W, int H, const float* matrix, const float* eye, float textureWeight, float time,
                             int maxBounces, float* accum, float* normal, float* position, int* hitId, float* depth,
                             int nthreads) {
    set_frame(matrix, eye, textureWeight, time, maxBounces);
    FrameRays fr = makeFrameRays();
    uint64_t ops = 0;
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 4) num_threads(nthreads) reduction(+ : ops)
    for (int y = 0; y < H; y++) {
#ifdef SAIL_COUNT_OPS
        g_sail_ops = 0;
#endif
        for (int x = 0; x < W; x++) {
            size_t i = (size_t)y * W + x;
            vec3 cache(accum[3 * i], accum[3 * i + 1], accum[3 * i + 2]);
            PixelOut o = fstrace_main(fr, x, y, W, H, cache, nullptr);
            accum[3 * i] = o.color.x.v; accum[3 * i + 1] = o.color.y.v; accum[3 * i + 2] = o.color.z.v;
            if (normal) { normal[3 * i] = o.normal.x.v; normal[3 * i + 1] = o.normal.y.v; normal[3 * i + 2] = o.normal.z.v; }
            if (position) { position[3 * i] = o.position.x.v; position[3 * i + 1] = o.position.y.v; position[3 * i + 2] = o.position.z.v; }
            if (hitId) hitId[i] = o.hitId;
            if (depth) depth[i] = o.depth.v;
        }
#ifdef SAIL_COUNT_OPS
        ops += g_sail_ops;
#endif
    }
    return ops;
}

// Per-bounce dump of one frame: hitId/t [W*H*maxBounces], rays [W*H*maxBounces*6]; radiance e [W*H*3]
// (the un-accumulated sample).  Unreached bounces: hitId = -2.
int oracle_trace_dump(int W, int H, const float* matrix, const float* eye, float time, int maxBounces, int* hitId,
                      float* t, float* rays, float* radiance) {
    set_frame(matrix, eye, 0.0f, time, maxBounces);
    FrameRays fr = makeFrameRays();
#pragma omp parallel for schedule(dynamic, 4)
    for (int y = 0; y < H; y++) {
        for (int x = 0; x < W; x++) {
            size_t i = (size_t)y * W + x;
            for (int b = 0; b < maxBounces; b++) {
                hitId[i * maxBounces + b] = -2;
                t[i * maxBounces + b] = 0.0f;
                for (int k = 0; k < 6; k++) rays[(i * maxBounces + b) * 6 + k] = 0.0f;
            }
            BounceDump d{hitId + i * maxBounces, t + i * maxBounces, rays + i * maxBounces * 6};
            PixelOut o = fstrace_main(fr, x, y, W, H, vec3(0.0f, 0.0f, 0.0f), &d);
            if (radiance) { radiance[3 * i] = o.color.x.v; radiance[3 * i + 1] = o.color.y.v; radiance[3 * i + 2] = o.color.z.v; }
        }
    }
    return 0;
}

// intersectObjects on arbitrary rays (fuzz / replay tests).  rays: N x 6 (origin, dir).
// out: id (-1 miss), t, normal N x 3 (after the flip of shader.shape.js:39-40), matCategory.
int oracle_intersect_rays(int N, const float* rays, int* id, float* t, float* normal, int* matCategory) {
#pragma omp parallel for schedule(static)
    for (int i = 0; i < N; i++) {
        Ray r(vec3(rays[6 * i], rays[6 * i + 1], rays[6 * i + 2]), vec3(rays[6 * i + 3], rays[6 * i + 4], rays[6 * i + 5]));
        Intersect ins = intersectObjects(r);
        bool miss = ins.d >= real(MAX_DISTANCE);
        id[i] = miss ? -1 : ins.index;
        t[i] = ins.d.v;
        if (normal) { normal[3 * i] = ins.normal.x.v; normal[3 * i + 1] = ins.normal.y.v; normal[3 * i + 2] = ins.normal.z.v; }
        if (matCategory) matCategory[i] = miss ? 0 : ins.matCategory;
    }
    return 0;
}

// Full hit record for unit KATs: out[24] = d, hit(3), normal(3), dpdu(3), sc(3), emission(3), into, matCategory, index.
int oracle_intersect_detail(const float* ray, float* out) {
    Ray r(vec3(ray[0], ray[1], ray[2]), vec3(ray[3], ray[4], ray[5]));
    Intersect ins = intersectObjects(r);
    bool miss = ins.d >= real(MAX_DISTANCE);
    memset(out, 0, 24 * sizeof(float));
    out[0] = ins.d.v;
    out[19] = miss ? -1.0f : (float)ins.index;
    if (miss) return 0;
    const vec3* v[5] = {&ins.hit, &ins.normal, &ins.dpdu, &ins.sc, &ins.emission};
    for (int k = 0; k < 5; k++) { out[1 + 3 * k] = v[k]->x.v; out[2 + 3 * k] = v[k]->y.v; out[3 + 3 * k] = v[k]->z.v; }
    out[16] = ins.into ? 1.0f : 0.0f;
    out[17] = (float)ins.matCategory;
    out[18] = ins.matIndex.v;
    return 0;
}

// random2(seed) at a pixel (random.glsl:5-10), for RNG KATs.
void oracle_random2(int x, int y, float seed, float* out) {
    gl_FragCoord = vec3((float)x + 0.5f, (float)y + 0.5f, 0.5f);
    vec2 u = random2(real(seed));
    out[0] = u.x.v;
    out[1] = u.y.v;
}

// Elementwise detmath KATs: which = 0 sin,1 cos,2 tan,3 atan,4 acos,5 asin,6 exp,7 log; 8 atan2(a,b); 9 pow(a,b)
void oracle_detmath(int which, int N, const float* a, const float* b, float* out) {
    for (int i = 0; i < N; i++) {
        float x = a[i], y = b ? b[i] : 0.0f, r = 0.0f;
        switch (which) {
            case 0: r = dm_sin(x); break;
            case 1: r = dm_cos(x); break;
            case 2: r = dm_tan(x); break;
            case 3: r = dm_atan(x); break;
            case 4: r = dm_acos(x); break;
            case 5: r = dm_asin(x); break;
            case 6: r = dm_exp(x); break;
            case 7: r = dm_log(x); break;
            case 8: r = dm_atan2(x, y); break;
            case 9: r = dm_pow(x, y); break;
            case 12: r = dm_rcp(x); break;
        }
        out[i] = r;
    }
}

// Single shader functions for known-answer tests.  `in`/`out` are small float arrays:
//   0 frDielectric(cosThetaI, etaI, etaT) -> 1          fresnel.glsl:31-46
//   1 frConductor(cosThetaI, etaI[3], etaT[3], k[3]) -> 3  fresnel.glsl:48-70
//   2 trowbridgeReitz_d(ax, ay, wh[3]) -> 1             microfacet.glsl:61-67
//   3 trowbridgeReitz_sample_wh(u[2], ax, ay, wo[3]) -> 3   microfacet.glsl:41-59
//   4 trowbridgeReitz_pdf(ax, ay, wo[3], wh[3]) -> 1    microfacet.glsl:69-71
//   5 getSurfaceColor(hit[3], uv[2], texparams row) -> 3 (scene set)  shader.texture.js:22-29
//   6 cosineSampleHemisphere(u[2]) -> 3, 7 uniformSampleSphere(u[2]) -> 3, 8 concentricSampleDisk(u[2]) -> 2   sampler.glsl
//   9 quadratic(A, B, C) -> ok, t0, t1                  utility.glsl:37-51
//  10 light arm of row `index`: in = pixel x, y, seed, hit[3], normal[3], index -> 3 (scene set)  area/point/spot.glsl
int oracle_unit(int which, const float* in, float* out) {
    auto v3 = [&](int o) { return vec3(in[o], in[o + 1], in[o + 2]); };
    auto put3 = [&](vec3 v) { out[0] = v.x.v; out[1] = v.y.v; out[2] = v.z.v; };
    switch (which) {
        case 0: out[0] = frDielectric(real(in[0]), real(in[1]), real(in[2])).v; return 0;
        case 1: put3(frConductor(real(in[0]), v3(1), v3(4), v3(7))); return 0;
        case 2: out[0] = trowbridgeReitz_d(real(in[0]), real(in[1]), v3(2)).v; return 0;
        case 3: put3(trowbridgeReitz_sample_wh(vec2(in[0], in[1]), real(in[2]), real(in[3]), v3(4))); return 0;
        case 4: out[0] = trowbridgeReitz_pdf(real(in[0]), real(in[1]), v3(2), v3(5)).v; return 0;
        case 5: put3(getSurfaceColor(v3(0), vec2(in[3], in[4]), rowCoord((int)in[5], tn_))); return 0;
        case 6: put3(cosineSampleHemisphere(vec2(in[0], in[1]))); return 0;
        case 7: put3(uniformSampleSphere(vec2(in[0], in[1]))); return 0;
        case 8: { vec2 d = concentricSampleDisk(vec2(in[0], in[1])); out[0] = d.x.v; out[1] = d.y.v; return 0; }
        case 9: { real t0 = 0.0f, t1 = 0.0f; bool ok = quadratic(real(in[0]), real(in[1]), real(in[2]), t0, t1);
                  out[0] = ok ? 1.0f : 0.0f; out[1] = t0.v; out[2] = t1.v; return 0; }
        case 10: {
            gl_FragCoord = vec3(in[0] + 0.5f, in[1] + 0.5f, 0.5f);
            int index = (int)in[9];
            if (index < 0 || index >= ln_) return -1;
            real row = rowCoord(index, ln_);
            int cat = readInt(lights, vec2(0.0f, row), LIGHTS_LENGTH);
            if (cat == AREA) put3(area_sample(parseArea(row), real(in[2]), v3(3), v3(6)));
            else if (cat == POINT) put3(point_sample(parsePoint(row), real(in[2]), v3(3), v3(6)));
            else if (cat == SPOT) put3(spot_sample(parseSpot(row), real(in[2]), v3(3), v3(6)));
            else return -1;
            return 0;
        }
    }
    return -1;
}

int oracle_counts_ops(void) {
#ifdef SAIL_COUNT_OPS
    return 1;
#else
    return 0;
#endif
}

}  // extern "C"
