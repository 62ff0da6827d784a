// sail_queue.cuh -- the wavefront tracer with its ray queues in device memory: one kernel per stage.
//
// This is the structure north_star sketches: camera-ray generation, nearest-hit intersection, shading (BSDF sampling, next-event
// estimation with its shadow ray, Russian roulette) and accumulation as SEPARATE kernels over path queues that are regrouped per
// hit object between the stages.  The shared-memory wavefront (sail_wavefront.cuh) keeps the queue of 512 paths inside one CTA and
// pays a CTA barrier pair per bounce -- 35 % of its stall samples on the quadric scene (profiles/r02_notes.md).  Here a batch is
// hundreds of thousands of paths, every stage is a grid of independent warps (no barrier anywhere), every stage has its own
// register budget, and the regrouping is global: after the counting sort a warp of the shade stage holds 32 paths that hit the
// same object (one shape code, one material, one texture).
//
//   batch  = npix pixels x nframes frames of one sail_render_n() call; path p = k * npix + pixel (k = frame within the call)
//   state  = three arrays, one record per path: ray (32 B: origin, direction), carry (32 B: throughput, radiance), hit (8 B: distance,
//            object or -1 = ended) -- split so that every stage reads whole 32-byte sectors of exactly what it needs
//   stages = raygen | per bounce: scan (+ histogram of hit objects) -> scatter (counting sort -> permutation) -> shade | accumulate
//
// Queue traffic per path and bounce: scan reads 40 B and writes 8 B, scatter reads 8 B and writes 4 B, shade reads 4 + 72 B and
// writes 64 (+ 8) B = 200 B.  A path is an independent computation (the hash RNG is keyed by pixel, frame time and bounce), so which thread
// evaluates it changes no bit: the output equals the megakernel's and the oracle's exactly.
#ifndef SAIL_QUEUE_CUH
#define SAIL_QUEUE_CUH
#include "sail_mega.cuh"

#define SAIL_Q_THREADS 128
#define SAIL_Q_MAX_BOUNCES 32
#define SAIL_Q_CLASSES (SAIL_MAX_OBJECTS + 1)  // one per object + "path has ended"

struct SailQLaunch {
    int W, H;
    int pix0, npix;      // this batch: pixels [pix0, pix0 + npix) in row-major order (row 0 = bottom); npix a multiple of 32
    int nframes, maxBounces, rrStart, accumMode;
    int bounce;          // 1-based bounce the scan / scatter / shade launch works on
    int pad0;
    float rcpW, rcpH;
    float* accum;                 // W*H*3
    const SailSceneBlock* scene;
    float4* ray;                  // [npix * nframes][2]: {o.xyz, d.x} {d.yz, -, -}
    float4* carry;                // [npix * nframes][2]: {fpdf.xyz, e.x} {e.yz, -, -}
    int2* hit;                    // [npix * nframes]: {float bits of t, object id or -1 = the path has ended}
    int* perm;                    // [npix * nframes] paths in shade order (regrouped by hit object)
    int* counters;                // [SAIL_Q_MAX_BOUNCES + 1][2][SAIL_Q_CLASSES]: histogram and scatter cursors per bounce
    SailFrameDev frames[SAIL_MAX_BATCH];
};

SD int* qCount(const SailQLaunch& L, int bounce) { return L.counters + (bounce * 2 + 0) * SAIL_Q_CLASSES; }
SD int* qCursor(const SailQLaunch& L, int bounce) { return L.counters + (bounce * 2 + 1) * SAIL_Q_CLASSES; }

// ---- stage 0: camera rays (vstrace.glsl:4-7 + the rasteriser's interpolation); also clears the batch's counters ------------------
template <int F>
SD void sailQRaygen(const SailQLaunch& L) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (blockIdx.x == 0)
        for (int i = threadIdx.x; i < (SAIL_Q_MAX_BOUNCES + 1) * 2 * SAIL_Q_CLASSES; i += blockDim.x) L.counters[i] = 0;
    if (p >= L.npix * L.nframes) return;
    const int k = p / L.npix, pix = L.pix0 + (p - k * L.npix);
    const int y = pix / L.W, x = pix - y * L.W;
    const SailFrameDev& f = L.frames[k];  // k is uniform across a warp (npix is a multiple of 32)
    const float fx = (float)x + 0.5f, fy = (float)y + 0.5f;
    const V3 d = primaryDir(f, fx * L.rcpW, fy * L.rcpH);
    L.ray[2 * (size_t)p] = make_float4(f.eye[0], f.eye[1], f.eye[2], d.x);
    L.ray[2 * (size_t)p + 1] = make_float4(d.y, d.z, 0.0f, 0.0f);
    L.carry[2 * (size_t)p] = make_float4(1.0f, 1.0f, 1.0f, 0.0f);
    L.carry[2 * (size_t)p + 1] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    L.hit[p] = make_int2(__float_as_int(SAIL_MAX_DISTANCE), 0);  // id >= 0: alive
}

// ---- stage 1: nearest-hit scan over the batch + histogram of the hit objects (shader.shape.js:28-37) --------------------------------
template <int F>
SD void sailQScan(const SailQLaunch& L) {
    __shared__ int hist[SAIL_Q_CLASSES];
    const SceneView sc = stageScene(L.scene);
    for (int i = threadIdx.x; i < SAIL_Q_CLASSES; i += blockDim.x) hist[i] = 0;
    __syncthreads();
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    const bool inRange = p < L.npix * L.nframes;
    int key = sc.n;  // ended
    if (inRange) {
        if (L.hit[p].y >= 0) {  // still alive
            const float4 s0 = L.ray[2 * (size_t)p], s1 = L.ray[2 * (size_t)p + 1];
            Ray ray;
            ray.o = mk3(s0.x, s0.y, s0.z);
            ray.d = mk3(s0.w, s1.x, s1.y);
            Nearest nr = nearestObject<F>(sc, ray);
            bool alive = nr.id >= 0;  // path.glsl:22 `break`
            // last bounce on an object that neither emits nor gets NEE: adds +0 to e (see tracePath()), the path is done
            if (alive && SAIL_IS_LAST_BOUNCE(L.bounce, L.maxBounces) && !lastBounceNeedsRecord<F>(sc.shape[nr.id])) alive = false;
            L.hit[p] = make_int2(__float_as_int(nr.t), alive ? nr.id : -1);
            if (alive) key = nr.id;
        }
    }
    // one shared-memory atomic per (warp, object), one global atomic per (CTA, object)
    const unsigned active = __ballot_sync(0xffffffffu, inRange);
    if (inRange) {
        const unsigned peers = __match_any_sync(active, key);
        if ((int)(threadIdx.x & 31) == __ffs(peers) - 1) atomicAdd(&hist[key], __popc(peers));
    }
    __syncthreads();
    int* count = qCount(L, L.bounce);
    for (int i = threadIdx.x; i <= sc.n; i += blockDim.x)
        if (hist[i]) atomicAdd(&count[i], hist[i]);
}

// ---- stage 2: counting sort -> permutation.  The order of the paths inside one object's range is arbitrary (and irrelevant). ---------
template <int F>
SD void sailQScatter(const SailQLaunch& L) {
    __shared__ int offset[SAIL_Q_CLASSES];
    const int n = L.scene->n;
    if (threadIdx.x == 0) {
        const int* count = qCount(L, L.bounce);
        int acc = 0;
        for (int i = 0; i <= n; i++) { offset[i] = acc; acc += count[i]; }
    }
    __syncthreads();
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    const bool inRange = p < L.npix * L.nframes;
    const unsigned active = __ballot_sync(0xffffffffu, inRange);
    if (!inRange) return;
    const int id = L.hit[p].y;
    const int key = id >= 0 ? id : n;
    const unsigned peers = __match_any_sync(active, key);
    const int lane = threadIdx.x & 31, leader = __ffs(peers) - 1;
    int base = 0;
    if (lane == leader) base = atomicAdd(&qCursor(L, L.bounce)[key], __popc(peers));
    base = __shfl_sync(peers, base, leader);
    L.perm[offset[key] + base + __popc(peers & ((1u << lane) - 1u))] = p;
}

// ---- stage 3: hit record + shading of the regrouped paths (path.glsl:21-36) --------------------------------------------------------
template <int F>
SD void sailQShade(const SailQLaunch& L) {
    const SceneView sc = stageScene(L.scene);
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= L.npix * L.nframes) return;
    // primary rays are coherent in pixel order already; later bounces run in the order of the counting sort
    const int p = L.bounce > 1 ? L.perm[j] : j;
    const int2 hit = L.hit[p];
    const int id = hit.y;
    if (id < 0) return;  // ended paths sort to the end: whole warps leave here
    const float4 s0 = L.ray[2 * (size_t)p], s1 = L.ray[2 * (size_t)p + 1];
    const float4 c0 = L.carry[2 * (size_t)p], c1 = L.carry[2 * (size_t)p + 1];
    Ray ray;
    ray.o = mk3(s0.x, s0.y, s0.z);
    ray.d = mk3(s0.w, s1.x, s1.y);
    V3 fpdf = mk3(c0.x, c0.y, c0.z), e = mk3(c0.w, c1.x, c1.y);
    const int k = p / L.npix, pix = L.pix0 + (p - k * L.npix);
    const int y = pix / L.W, x = pix - y * L.W;
    const float fx = (float)x + 0.5f, fy = (float)y + 0.5f;
    Hit h;
    hitRecord<F>(sc, id, __int_as_float(hit.x), ray, h);
    bool alive = true;
    if (SAIL_IS_LAST_BOUNCE(L.bounce, L.maxBounces) && lastBounceIsEmissionOnly<F>(h)) {  // see tracePath(): dead BSDF sample
        e = e + (h.emission + splat3(0.0f)) * fpdf;
        alive = false;
    } else {
        // after the sort k differs from lane to lane: the constant-bank read of the frame time serialises per distinct k, once per bounce
        const float seed = L.frames[k].timeSinceStart + (float)L.bounce;
        V2 u = random2<F>(fx, fy, 0.5f, seed);
        V3 wi, _fpdf;
        e = e + shade<F>(sc, h, u, -ray.d, wi, _fpdf) * fpdf;
        fpdf = fpdf * _fpdf;
        if (!russianRoulette<F>(L.rrStart, L.bounce, L.maxBounces, fx, fy, 0.5f, seed, fpdf)) alive = false;
        const float outdot = dot3(h.n, wi);
        ray.o = h.p + h.n * (outdot > SAIL_EPSILON ? 0.0001f : -0.0001f);
        ray.d = wi;
    }
    if (L.bounce >= L.maxBounces) alive = false;
    if (alive) {
        L.ray[2 * (size_t)p] = make_float4(ray.o.x, ray.o.y, ray.o.z, ray.d.x);
        L.ray[2 * (size_t)p + 1] = make_float4(ray.d.y, ray.d.z, 0.0f, 0.0f);
    } else {
        L.hit[p] = make_int2(hit.x, -1);
    }
    L.carry[2 * (size_t)p] = make_float4(fpdf.x, fpdf.y, fpdf.z, e.x);
    L.carry[2 * (size_t)p + 1] = make_float4(e.y, e.z, 0.0f, 0.0f);
}

// ---- stage 4: progressive accumulation of the batch's frames, in frame order (fstrace.glsl:13-14) ---------------------------------
template <int F>
SD void sailQAccumulate(const SailQLaunch& L) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= L.npix) return;
    const size_t pix = (size_t)L.pix0 + i;
    V3 c = mk3(L.accum[3 * pix], L.accum[3 * pix + 1], L.accum[3 * pix + 2]);
    for (int k = 0; k < L.nframes; k++) {
        const size_t p = (size_t)k * L.npix + i;
        const float4 c0 = L.carry[2 * p], c1 = L.carry[2 * p + 1];
        const V3 e = mk3(nanToZero(c0.w), nanToZero(c1.x), nanToZero(c1.y));
        if (L.accumMode == 0) c = mix3(e, c, L.frames[k].textureWeight);
        else c = c + e;
    }
    L.accum[3 * pix] = c.x;
    L.accum[3 * pix + 1] = c.y;
    L.accum[3 * pix + 2] = c.z;
}

#endif
