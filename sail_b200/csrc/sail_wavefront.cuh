// sail_wavefront.cuh -- the wavefront variant, with its ray queues in SHARED memory.
//
// north_star sketches a wavefront tracer: per-bounce stages over ray queues compacted per material.  With
// queues in HBM these scenes (a few hundred bytes of geometry, all path state in registers) would become
// bandwidth bound: ~60 B of state written and re-read per path per bounce is 4-5 TB/s at the megakernel's
// rate.  On B200 a CTA has up to 227 KB of shared memory, so the queue lives there instead:
// each CTA is a persistent 128-path wavefront that, every bounce,
//   1. runs the nearest-hit scan for its paths (uniform code, no divergence),
//   2. counting-sorts the 128 paths by the object they hit (shared-memory histogram + prefix),
//   3. exchanges path state through shared memory so that consecutive lanes hold paths that hit the same
//      object -- hence the same shape code, material and texture,
//   4. evaluates the hit record and the BSDF / light sampling on the regrouped paths.
// A path is an independent computation, so regrouping changes which lane computes it but not one bit of its
// value: the output equals the megakernel's exactly (tests/test_gpu_parity.py).  Paths carry their home lane
// and return their radiance sample to it at the end of each frame; the home lane owns the accumulator.
#ifndef SAIL_WAVEFRONT_CUH
#define SAIL_WAVEFRONT_CUH
#include "sail_mega.cuh"

#define SAIL_WF_STATE 16  // words of path state exchanged per bounce
// Paths per wavefront (= threads per CTA).  128 in the ahead-of-time variants; the run-time compiled kernel may use more
// (sail_jit.cpp): a larger wavefront regroups over more paths (fuller warps per object) and keeps more of an SM's warps in
// the same phase of the bounce, which matters once the kernel is larger than the 32 KB instruction cache.  The CTA's pixel
// tile is (SAIL_WF_THREADS / 8) x 8: two rows of 8x4 warp sub-tiles, so row bands keep their 8-row granularity.
#ifndef SAIL_WF_THREADS
#define SAIL_WF_THREADS SAIL_CTA_THREADS
#endif
// Bounces before which the paths are regrouped (bit d = bounce d; A/B knob, default: every bounce after the first).
#ifndef SAIL_WF_REGROUP_MASK
#define SAIL_WF_REGROUP_MASK 0xffffffffu
#endif
#define SAIL_WF_TILE_W (SAIL_WF_THREADS / 8)
#define SAIL_WF_WARPS_X (SAIL_WF_TILE_W / 8)

template <int F>
SD void sailWaveBody(const SailLaunch& L) {
    __shared__ float slot[SAIL_WF_STATE][SAIL_WF_THREADS];  // SoA: lane-consecutive words, conflict free
    __shared__ int hist[2][SAIL_MAX_OBJECTS + 2];  // double buffered: the next bounce's counters are cleared while this one scatters
    const SceneView sc = stageScene(L.scene);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int x = blockIdx.x * SAIL_WF_TILE_W + (warp % SAIL_WF_WARPS_X) * 8 + (lane & 7);
    int tileY = L.y0 + blockIdx.y * SAIL_TILE_H;
    if (L.bandStride > 1) {
        const int tpb = L.bandRows / SAIL_TILE_H;
        tileY = (L.bandFirst + ((int)blockIdx.y / tpb) * L.bandStride) * L.bandRows + ((int)blockIdx.y % tpb) * SAIL_TILE_H;
    }
    const int y = tileY + (warp / SAIL_WF_WARPS_X) * 4 + (lane >> 3);
    const bool inside = x < L.W && y < L.y1;  // no early exit: every lane takes part in the barriers
    const size_t pix = inside ? (size_t)y * L.W + x : 0;
    const float hx = (float)x + 0.5f, hy = (float)y + 0.5f;  // gl_FragCoord of the home pixel
    const float a = hx * L.rcpW, b = hy * L.rcpH;
    const int n = sc.n;
    if (tid <= n) { hist[0][tid] = 0; hist[1][tid] = 0; }
    int hsel = 0;  // which histogram the next regroup uses (uniform across the CTA); stageScene's barrier is behind us,
    __syncthreads();  // this one publishes the cleared counters

    V3 c = inside ? mk3(L.accum[3 * pix], L.accum[3 * pix + 1], L.accum[3 * pix + 2]) : splat3(0.0f);

    for (int k = 0; k < L.nframes; k++) {
        const SailFrameDev& f = L.frames[k];
        // path state (travels between lanes)
        Ray ray;
        ray.o = ld3(f.eye);
        ray.d = primaryDir(f, a, b);
        V3 fpdf = splat3(1.0f), e = splat3(0.0f);
        float fx = hx, fy = hy;  // RNG key of the path's pixel
        int home = tid;
        bool alive = inside;

        for (int deepth = 1; deepth <= L.maxBounces; deepth++) {
            // ---- stage 1: nearest-hit scan (uniform) ----
            Nearest nr;
            nr.t = SAIL_MAX_DISTANCE;
            nr.id = -1;
            if (alive) nr = nearestObject<F>(sc, ray);
            if (nr.id < 0) alive = false;                          // path.glsl:22 `break`
            // last bounce on an object that neither emits nor gets NEE: adds +0 to e (see tracePath()), the path is done
            if (alive && SAIL_IS_LAST_BOUNCE(deepth, L.maxBounces) && !lastBounceNeedsRecord<F>(sc.shape[nr.id])) alive = false;
            float t = nr.t;
            int id = alive ? nr.id : -1;
            // Primary rays of a tile are coherent already; regroup from the second bounce on.
            if (deepth > 1 && ((SAIL_WF_REGROUP_MASK >> (deepth & 31)) & 1u)) {
                // ---- stage 2: counting sort by hit object ----
                const int key = alive ? nr.id : n;
                int* hcur = hist[hsel];
                // one shared-memory atomic per (warp, object) instead of one per lane: lanes that hit the same object find each
                // other with match.any, the first of them reserves the whole group's ranks (same-address atomics of a warp are
                // serialised by the hardware -- 4.3e8 bank conflicts per launch in the round-1 capture)
                const unsigned peers = __match_any_sync(0xffffffffu, key);
                const int leader = __ffs(peers) - 1;
                int base = 0;
                if (lane == leader) base = atomicAdd(&hcur[key], __popc(peers));
                base = __shfl_sync(0xffffffffu, base, leader);
                const int rank = base + __popc(peers & ((1u << lane) - 1u));
                __syncthreads();
                int pos = rank;
                for (int j = 0; j < key; j++) pos += hcur[j];
                if (tid <= n) hist[hsel ^ 1][tid] = 0;  // nobody reads or counts in the other buffer before the next barrier
                hsel ^= 1;
                // ---- stage 3: exchange ----
                slot[0][pos] = ray.o.x; slot[1][pos] = ray.o.y; slot[2][pos] = ray.o.z;
                slot[3][pos] = ray.d.x; slot[4][pos] = ray.d.y; slot[5][pos] = ray.d.z;
                slot[6][pos] = fpdf.x; slot[7][pos] = fpdf.y; slot[8][pos] = fpdf.z;
                slot[9][pos] = e.x; slot[10][pos] = e.y; slot[11][pos] = e.z;
                slot[12][pos] = fx; slot[13][pos] = fy;
                slot[14][pos] = nr.t;
                slot[15][pos] = __int_as_float((home << 8) | (alive ? (nr.id + 1) : 0));
                __syncthreads();
                ray.o = mk3(slot[0][tid], slot[1][tid], slot[2][tid]);
                ray.d = mk3(slot[3][tid], slot[4][tid], slot[5][tid]);
                fpdf = mk3(slot[6][tid], slot[7][tid], slot[8][tid]);
                e = mk3(slot[9][tid], slot[10][tid], slot[11][tid]);
                fx = slot[12][tid]; fy = slot[13][tid];
                t = slot[14][tid];
                const int tag = __float_as_int(slot[15][tid]);
                home = tag >> 8;
                id = (tag & 0xff) - 1;
                alive = id >= 0;
            }
            // ---- stage 4: hit record + shading on regrouped paths (path.glsl:21-36) ----
            if (alive) {
                Hit h;
                hitRecord<F>(sc, id, t, ray, h);
                if (SAIL_IS_LAST_BOUNCE(deepth, L.maxBounces) && lastBounceIsEmissionOnly<F>(h)) {  // see tracePath(): dead BSDF sample
                    e = e + (h.emission + splat3(0.0f)) * fpdf;
                } else {
                    const float seed = f.timeSinceStart + (float)deepth;
                    V2 u = random2<F>(fx, fy, 0.5f, seed);
                    V3 wi, _fpdf;
                    e = e + shade<F>(sc, h, u, -ray.d, wi, _fpdf) * fpdf;
                    fpdf = fpdf * _fpdf;
                    if (!russianRoulette<F>(L.rrStart, deepth, L.maxBounces, fx, fy, 0.5f, seed, fpdf)) alive = false;
                    const float outdot = dot3(h.n, wi);
                    ray.o = h.p + h.n * (outdot > SAIL_EPSILON ? 0.0001f : -0.0001f);
                    ray.d = wi;
                }
            }
        }
        // ---- return the sample to its home lane ----
        __syncthreads();
        slot[0][home] = e.x; slot[1][home] = e.y; slot[2][home] = e.z;
        __syncthreads();
        V3 es = mk3(nanToZero(slot[0][tid]), nanToZero(slot[1][tid]), nanToZero(slot[2][tid]));
        if (L.accumMode == 0) c = mix3(es, c, f.textureWeight);  // fstrace.glsl:14
        else c = c + es;
        if (k + 1 < L.nframes) __syncthreads();  // slot[] is rewritten by the next frame
    }
    if (!inside) return;
    L.accum[3 * pix] = c.x;
    L.accum[3 * pix + 1] = c.y;
    L.accum[3 * pix + 2] = c.z;
}

template <int F>
__global__ void __launch_bounds__(SAIL_CTA_THREADS, SAIL_MIN_CTAS(F)) sail_trace_wavefront(const __grid_constant__ SailLaunch L) {
    sailWaveBody<F>(L);
}

#endif
