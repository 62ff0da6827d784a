// sail_diag.cuh -- diagnostic / picking kernels behind sail_intersect_rays(), sail_trace_dump(),
// sail_eval_math() and sail_scale_accum() (include/sail_b200.h).  Built with every feature enabled.
#ifndef SAIL_DIAG_CUH
#define SAIL_DIAG_CUH
#include "sail_mega.cuh"

// ---- diagnostics ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(SAIL_CTA_THREADS) sail_intersect_kernel(const SailSceneBlock* __restrict__ scene,
                                                                          const float* __restrict__ rays, int count, int* hitId,
                                                                          float* t, float* normal, int* matCategory) {
    const SceneView sc = stageScene(scene);
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    Ray r;
    r.o = ld3(rays + 6 * i);
    r.d = ld3(rays + 6 * i + 3);
    Hit h;
    h.n = splat3(0.0f);
    h.matCategory = 0;
    bool ok = intersectObjects<SAIL_F_ALL>(sc, r, h);
    if (hitId) hitId[i] = h.index;
    if (t) t[i] = h.d;
    if (normal) { normal[3 * i] = ok ? h.n.x : 0.0f; normal[3 * i + 1] = ok ? h.n.y : 0.0f; normal[3 * i + 2] = ok ? h.n.z : 0.0f; }
    if (matCategory) matCategory[i] = ok ? h.matCategory : 0;
}

struct SailDumpLaunch {
    int W, H, maxBounces, rrStart;
    float rcpW, rcpH;
    const SailSceneBlock* scene;
    int* hitId;
    float* t;
    float* rays;
    float* radiance;
    SailFrameDev frame;
};
// trace() with per-bounce records; the loop body is the one of tracePath().
__global__ void __launch_bounds__(SAIL_CTA_THREADS) sail_trace_dump_kernel(const __grid_constant__ SailDumpLaunch L) {
    const SceneView sc = stageScene(L.scene);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int x = blockIdx.x * SAIL_TILE_W + (warp & 1) * 8 + (lane & 7);
    const int y = blockIdx.y * SAIL_TILE_H + (warp >> 1) * 4 + (lane >> 3);
    if (x >= L.W || y >= L.H) return;
    const size_t pix = (size_t)y * L.W + x;
    const float fx = (float)x + 0.5f, fy = (float)y + 0.5f, fz = 0.5f;
    Ray ray;
    ray.o = ld3(L.frame.eye);
    ray.d = primaryDir(L.frame, fx * L.rcpW, fy * L.rcpH);
    const int B = L.maxBounces;
    for (int k = 0; k < B; k++) {
        if (L.hitId) L.hitId[pix * B + k] = -2;
        if (L.t) L.t[pix * B + k] = 0.0f;
        if (L.rays) for (int j = 0; j < 6; j++) L.rays[(pix * B + k) * 6 + j] = 0.0f;
    }
    V3 fpdf = splat3(1.0f), e = splat3(0.0f);
    int deepth = 0;
    while (deepth++ < B) {
        Hit h;
        bool ok = intersectObjects<SAIL_F_ALL>(sc, ray, h);
        size_t o = pix * B + (deepth - 1);
        if (L.hitId) L.hitId[o] = h.index;
        if (L.t) L.t[o] = h.d;
        if (L.rays) {
            float* r = L.rays + o * 6;
            r[0] = ray.o.x; r[1] = ray.o.y; r[2] = ray.o.z; r[3] = ray.d.x; r[4] = ray.d.y; r[5] = ray.d.z;
        }
        if (!ok) break;
        float seed = L.frame.timeSinceStart + (float)deepth;
        V2 u = random2<SAIL_F_ALL>(fx, fy, fz, seed);
        V3 wi, _fpdf;
        e = e + shade<SAIL_F_ALL>(sc, h, u, -ray.d, wi, _fpdf) * fpdf;
        fpdf = fpdf * _fpdf;
        if (!russianRoulette<SAIL_F_ALL>(L.rrStart, deepth, B, fx, fy, fz, seed, fpdf)) break;
        float outdot = dot3(h.n, wi);
        ray.o = h.p + h.n * (outdot > SAIL_EPSILON ? 0.0001f : -0.0001f);
        ray.d = wi;
    }
    if (L.radiance) {
        L.radiance[3 * pix] = nanToZero(e.x);
        L.radiance[3 * pix + 1] = nanToZero(e.y);
        L.radiance[3 * pix + 2] = nanToZero(e.z);
    }
}

__global__ void sail_eval_math_kernel(int which, const float* a, const float* b, float* out, int count) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
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
        case 10: r = dm_div(x, y); break;
        case 11: r = dm_sqrt(x); break;
        case 12: r = dm_rcp(x); break;
    }
    out[i] = r;
}

// Exhaustive check of dm_rcp over all 2^32 inputs against the rule stated in detmath.h (IEEE __frcp_rn inside
// 2^-126 <= |a| <= 2^126, pinned values outside).  One thread per 2^12 consecutive bit patterns.
__global__ void sail_selftest_rcp_kernel(unsigned long long* mismatches) {
    unsigned base = (blockIdx.x * blockDim.x + threadIdx.x) << 12;
    unsigned bad = 0;
    for (unsigned k = 0; k < 4096; k++) {
        unsigned u = base + k, mag = u & 0x7fffffffu, sign = u & 0x80000000u;
        float a = __uint_as_float(u), got = dm_rcp(a), want;
        if (mag > 0x7f800000u) want = a;
        else if (mag < 0x00800000u) want = __uint_as_float(sign | 0x7f800000u);
        else if (mag > 0x7e800000u) want = __uint_as_float(sign);
        else want = __frcp_rn(a);
        bool same = (got != got && want != want) || __float_as_uint(got) == __float_as_uint(want);
        bad += same ? 0u : 1u;
    }
    if (bad) atomicAdd(mismatches, (unsigned long long)bad);
}

// Exhaustive check of the domain-restricted helpers of the cosine sampler: sqrtUnit against sqrt.rn for x = 0 and every
// float in [2^-100, FLT_MAX] (plus NaN -> NaN), sincosUnit against dm_sincos for every float in [0, 8] and its negative.
// One thread per 2^12 consecutive positive bit patterns.
__global__ void sail_selftest_unit_kernel(unsigned long long* mismatches) {
    unsigned base = (blockIdx.x * blockDim.x + threadIdx.x) << 12;
    unsigned bad = 0;
    for (unsigned k = 0; k < 4096; k++) {
        const unsigned u = base + k;
        const float a = __uint_as_float(u);
        if (u == 0u || (u >= 0x0d800000u && u <= 0x7f7fffffu) || u > 0x7f800000u) {
            const float got = sqrtUnit(a), want = __fsqrt_rn(a);
            bad += ((got != got && want != want) || __float_as_uint(got) == __float_as_uint(want)) ? 0u : 1u;
        }
        if (u <= 0x41000000u) {  // [0, 8]
            for (int sgn = 0; sgn < 2; sgn++) {
                const float x = sgn ? -a : a;
                float s, c;
                dm_sincos(x, &s, &c);
                const float2 g = sincosUnit(x);
                bad += (__float_as_uint(g.x) == __float_as_uint(s) && __float_as_uint(g.y) == __float_as_uint(c)) ? 0u : 1u;
            }
        }
    }
    if (bad) atomicAdd(mismatches, (unsigned long long)bad);
}

__global__ void sail_scale_kernel(float* p, size_t count, float s) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) p[i] = p[i] * s;
}


#endif
