// sail_mega.cuh -- the megakernel: one thread carries one pixel through ray generation, the bounce
// loop, next-event estimation and progressive accumulation, for up to SAIL_MAX_FRAMES_PER_CALL
// consecutive frames per launch.
//
// Replaces the full-screen-quad draw of the trace program
// (/root/reference/src/core/webgl.js:82-83 -> src/shader/main/vstrace.glsl:4-7, main/fstrace.glsl:6-17)
// and the ping-pong colour accumulator (webgl.js:65-73,88-92).  B200 mapping:
//   * CTA = 128 threads = a 16x8 pixel tile, each warp an 8x4 sub-tile, so neighbouring lanes hit the
//     same objects / materials on the first bounces and the accumulator rows of a warp are contiguous;
//   * the packed scene block (sail_scene.h) is staged once per CTA into shared memory and read with
//     warp-uniform addresses in the object scan (broadcast, no bank conflicts);
//   * frame uniforms travel as a __grid_constant__ kernel parameter (constant bank, broadcast);
//   * the running mean of a pixel lives in registers across the frames of a launch: HBM traffic is one
//     12 B read + one 12 B write per pixel per launch instead of per frame (the reference re-reads and
//     re-writes its accumulation texture every frame).
#ifndef SAIL_MEGA_CUH
#define SAIL_MEGA_CUH
#include "sail_shade.cuh"

#define SAIL_MAX_BATCH 32
#define SAIL_TILE_W 16
#define SAIL_TILE_H 8
#define SAIL_CTA_THREADS 128
// Resident CTAs per SM the register allocation is capped for: 8 (64 registers, 32 warps/SM) for the small
// variants, 4 for those carrying microfacet BSDFs or quadrics (measured: profiles/r01_notes.md).
#ifndef SAIL_MIN_CTAS_SMALL
#define SAIL_MIN_CTAS_SMALL 8
#endif
#ifndef SAIL_MIN_CTAS_BIG
#define SAIL_MIN_CTAS_BIG 8
#endif
#define SAIL_MIN_CTAS(F) (((F) & (SAIL_F_GLOSSY | SAIL_F_QUADRICS)) ? SAIL_MIN_CTAS_BIG : SAIL_MIN_CTAS_SMALL)

// Per-frame constants derived on the host from (matrix, eye): the four corner rays of vstrace.glsl:4-7
// and the edge differences used by the rasteriser-style interpolation (DESIGN.md "Primary rays").
struct SailFrameDev {
    float d0[3], e20[3], e10[3];  // lower-left triangle: d0 + a*(d2-d0) + b*(d1-d0)
    float d3[3], e13[3], e23[3];  // upper-right triangle: d3 + (1-a)*(d1-d3) + (1-b)*(d2-d3)
    float eye[3];
    float textureWeight, timeSinceStart;
    float pad;
};

struct SailLaunch {
    int W, H;
    int y0, y1;          // rows rendered by this launch
    int bandRows, bandStride, bandFirst;  // interleaved row bands (bandStride > 1), see sail_set_row_bands()
    int maxBounces;
    int nframes;
    int accumMode;       // SAIL_ACCUM_*
    int rrStart;         // Russian roulette from this bounce on; 0 = off (sail_config.rr_start_bounce)
    float rcpW, rcpH;
    float* accum;        // W*H*3
    const SailSceneBlock* scene;
    unsigned sceneSmem;  // dynamic shared memory bytes (sailSceneSmemBytes)
    SailFrameDev frames[SAIL_MAX_BATCH];
};

// Stage the used part of the scene block into dynamic shared memory (16-byte copies) and return the view.
extern __shared__ uint4 sail_dyn_smem[];
SD SceneView stageScene(const SailSceneBlock* __restrict__ src) {
    const int n = src->n, ln = src->ln;
    const uint4* s4 = reinterpret_cast<const uint4*>(src);
    uint4* d4 = sail_dyn_smem;
    const int tid = threadIdx.x, nt = blockDim.x;
    const int shapeOff = SAIL_SCENE_SHAPE_OFF / 16, surfOff = SAIL_SCENE_SURF_OFF / 16, lightOff = SAIL_SCENE_LIGHT_OFF / 16;
    const int nShape = n * (int)(sizeof(SailShape) / 16), nSurf = n * (int)(sizeof(SailSurface) / 16);
    const int nLight = (ln + 1) * (int)(sizeof(SailLight) / 16);
    if (tid == 0) d4[0] = s4[0];
#pragma unroll 1
    for (int i = tid; i < nShape; i += nt) d4[1 + i] = s4[shapeOff + i];
#pragma unroll 1
    for (int i = tid; i < nSurf; i += nt) d4[1 + nShape + i] = s4[surfOff + i];
#pragma unroll 1
    for (int i = tid; i < nLight; i += nt) d4[1 + nShape + nSurf + i] = s4[lightOff + i];
    __syncthreads();
#ifndef SAIL_NO_OPAQUE_SMEM_BASE
    // Keep the shared-memory address of the block in ONE register.  Left to itself ptxas re-derives the window base
    // (S2UR SR_CgaCtaId + UMOV + ULEA) in front of every group of scene loads, 4 x 3 issue slots per bounce (measured:
    // Cornell +0.4 %, config 2 +1.8 %, config 5 +2.2 %).
    unsigned sbase = (unsigned)__cvta_generic_to_shared(sail_dyn_smem);
    asm volatile("mov.u32 %0, %0;" : "+r"(sbase));
    return sceneViewAt(reinterpret_cast<const unsigned char*>(__cvta_shared_to_generic(sbase)));
#else
    return sceneViewAt(reinterpret_cast<const unsigned char*>(sail_dyn_smem));
#endif
}
#ifndef __CUDACC_RTC__
// Dynamic shared memory a launch needs for a scene with n objects and ln lights.
static inline size_t sailSceneSmemBytes(int n, int ln) {
    return 16 + (size_t)n * (sizeof(SailShape) + sizeof(SailSurface)) + (size_t)(ln + 1) * sizeof(SailLight);
}
#endif

// Interpolated primary ray direction at the centre of pixel (x, y): not renormalised.
SD V3 primaryDir(const SailFrameDev& f, float a, float b) {
    // Both triangles evaluate base + s*e1 + t*e2: pick the operands, not the result (d0,e20,e10 and d3,e13,e23
    // are laid out back to back in SailFrameDev).
    const bool lower = a + b < 1.0f;
    const int o = lower ? 0 : 9;  // floats from d0 to d3
    const float s = lower ? a : 1.0f - a, t = lower ? b : 1.0f - b;
    return ld3(f.d0 + o) + s * ld3(f.d0 + o + 3) + t * ld3(f.d0 + o + 6);
}
SD float nanToZero(float v) { return (v == v) ? v : 0.0f; }

// The kernel body is a device function so that the ahead-of-time variants below and the scene-specialised kernels
// compiled at run time (sail_jit.cpp) share it; L is the kernel's __grid_constant__ parameter (constant bank).
template <int F>
SD void sailMegaBody(const SailLaunch& L) {
    const SceneView sc = stageScene(L.scene);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int x = blockIdx.x * SAIL_TILE_W + (warp & 1) * 8 + (lane & 7);
    int tileY = L.y0 + blockIdx.y * SAIL_TILE_H;
    if (L.bandStride > 1) {
        const int tpb = L.bandRows / SAIL_TILE_H;
        tileY = (L.bandFirst + ((int)blockIdx.y / tpb) * L.bandStride) * L.bandRows + ((int)blockIdx.y % tpb) * SAIL_TILE_H;
    }
    const int y = tileY + (warp >> 1) * 4 + (lane >> 3);
    if (x >= L.W || y >= L.y1) return;

    const size_t pix = (size_t)y * L.W + x;
    // gl_FragCoord of the full-screen quad at z = 0 (depth 0.5)
    float fx = (float)x + 0.5f, fy = (float)y + 0.5f;
    const float fz = 0.5f;
#ifndef SAIL_REMAT_FRAGCOORD
    // Opaque to the compiler: otherwise ptxas re-derives fx from %tid / %ctaid in every bounce (9 issue slots) instead of
    // keeping one register (or one local-memory word) alive.
    asm volatile("" : "+f"(fx), "+f"(fy));
#endif
    const float a = fx * L.rcpW, b = fy * L.rcpH;

    V3 c = mk3(L.accum[3 * pix], L.accum[3 * pix + 1], L.accum[3 * pix + 2]);
    for (int k = 0; k < L.nframes; k++) {
        const SailFrameDev& f = L.frames[k];
        Ray ray;
        ray.o = ld3(f.eye);
        ray.d = primaryDir(f, a, b);
        float tss = f.timeSinceStart;
#ifndef SAIL_RELOAD_TIME
        asm volatile("" : "+f"(tss));  // one register instead of three uniform-path instructions per bounce to re-read it
#endif
        V3 e = tracePath<F>(sc, ray, L.maxBounces, fx, fy, fz, tss, L.rrStart);
        e = mk3(nanToZero(e.x), nanToZero(e.y), nanToZero(e.z));
        if (L.accumMode == 0) c = mix3(e, c, f.textureWeight);  // fstrace.glsl:14
        else c = c + e;
    }
    L.accum[3 * pix] = c.x;
    L.accum[3 * pix + 1] = c.y;
    L.accum[3 * pix + 2] = c.z;
}

template <int F>
__global__ void __launch_bounds__(SAIL_CTA_THREADS, SAIL_MIN_CTAS(F)) sail_trace_megakernel(const __grid_constant__ SailLaunch L) {
    sailMegaBody<F>(L);
}

#endif
