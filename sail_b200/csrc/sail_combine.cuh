// sail_combine.cuh -- the one exchange step of the path: forming the image of a sharded (multi-GPU) context on its root.
//
// The reference has a single GL context and a single accumulation texture (/root/reference/src/core/webgl.js:12-28,215-227).
// A sharded context holds one FP32 buffer per GPU; when an image is asked for they are combined on the root:
//   * sample sharding: every member owns 1/N of the pixels; it reads that slice of ALL members' sums through NVLink peer
//     mappings (plain loads), adds them in member order 0..N-1 -- a fixed order, so the result is deterministic and equal
//     to a host-side sum in that order -- scales by 1/samples and stores the slice straight into the root's combined buffer
//     (peer stores).  Ingress per GPU is (N-1)/N of the image, spread over all NVSwitch ports, instead of (N-1) images into
//     the root.  This is the reduce-scatter + gather of a ring reduce fused into one kernel per GPU with no staging buffer.
//   * tile sharding: members copy their row bands into the root's buffer (strided peer copy, sail_api.cu); no arithmetic.
// Also here: the primary-hit AOV pass a sharded context runs on its root (the AOVs are those of the last frame,
// src/shader/main/fstrace.glsl:15-16, whichever member traced it).
#ifndef SAIL_COMBINE_CUH
#define SAIL_COMBINE_CUH
#include "sail_mega.cuh"

#define SAIL_COMBINE_MAX 8  // = SAIL_MAX_DEVICES of include/sail_b200.h

struct SailCombineParams {
    const float* src[SAIL_COMBINE_MAX];  // the members' sum buffers, member order
    float* dst;                          // the root's combined buffer
    int nsrc;
    unsigned long long begin, end;       // this member's slice in floats; begin is a multiple of 4
    float scale;                         // 1 / samples (SAIL_ACCUM_MEAN) or 1 (SAIL_ACCUM_SUM)
};

// HBM / NVLink bound: per float 4*nsrc bytes in, 4 bytes out.  One float4 per source per thread and iteration, all nsrc
// loads of an iteration in flight together.
__global__ void __launch_bounds__(256) sail_combine_sum_kernel(const __grid_constant__ SailCombineParams P) {
    const unsigned long long nvec = (P.end - P.begin) >> 2;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long v = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += stride) {
        const unsigned long long i = P.begin + 4ull * v;
        float4 a[SAIL_COMBINE_MAX];
#pragma unroll
        for (int r = 0; r < SAIL_COMBINE_MAX; r++)
            if (r < P.nsrc) a[r] = *reinterpret_cast<const float4*>(P.src[r] + i);
        float4 s = a[0];
#pragma unroll
        for (int r = 1; r < SAIL_COMBINE_MAX; r++)
            if (r < P.nsrc) { s.x = s.x + a[r].x; s.y = s.y + a[r].y; s.z = s.z + a[r].z; s.w = s.w + a[r].w; }
        s.x = s.x * P.scale; s.y = s.y * P.scale; s.z = s.z * P.scale; s.w = s.w * P.scale;
        *reinterpret_cast<float4*>(P.dst + i) = s;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0)  // the last slice may end off a float4 boundary
        for (unsigned long long i = P.begin + 4ull * nvec; i < P.end; i++) {
            float s = P.src[0][i];
            for (int r = 1; r < P.nsrc; r++) s = s + P.src[r][i];
            P.dst[i] = s * P.scale;
        }
}

struct SailAovLaunch {
    int W, H, maxBounces;
    float rcpW, rcpH;
    const SailSceneBlock* scene;
    float* normal;
    float* position;
    int* hitId;
    float* depth;
    SailFrameDev frame;
};
// The first-hit outputs of one frame without tracing it: exactly what tracePath() hands to writeAov() on bounce 1.
__global__ void __launch_bounds__(SAIL_CTA_THREADS) sail_aov_kernel(const __grid_constant__ SailAovLaunch L) {
    const SceneView sc = stageScene(L.scene);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int x = blockIdx.x * SAIL_TILE_W + (warp & 1) * 8 + (lane & 7);
    const int y = blockIdx.y * SAIL_TILE_H + (warp >> 1) * 4 + (lane >> 3);
    if (x >= L.W || y >= L.H) return;
    AovSink aov;
    aov.normal = L.normal; aov.position = L.position; aov.hitId = L.hitId; aov.depth = L.depth;
    aov.pix = (size_t)y * L.W + x;
    aov.on = true;
    if (L.maxBounces < 1) { writeAov(aov, splat3(0.0f), splat3(0.0f), -1, SAIL_MAX_DISTANCE); return; }
    const float fx = (float)x + 0.5f, fy = (float)y + 0.5f;
    Ray ray;
    ray.o = ld3(L.frame.eye);
    ray.d = primaryDir(L.frame, fx * L.rcpW, fy * L.rcpH);
    const Nearest nr = nearestObject<SAIL_F_ALL>(sc, ray);
    if (nr.id < 0) { writeAov(aov, splat3(0.0f), splat3(0.0f), nr.id, nr.t); return; }
    Hit h;
    hitRecord<SAIL_F_ALL>(sc, nr.id, nr.t, ray, h);
    writeAov(aov, h.n, h.p, h.index, h.d);
}

#endif
