// sail_server.cuh -- the megakernel with a *server warp*: warp specialisation against shape divergence.
//
// ncu on the megakernel (profiles/r01m_*): on the Cornell box 28 % of all issued warp instructions run with fewer
// than 8 of 32 lanes -- the sphere's hit record and mirror BSDF.  After the first bounce ~5 % of the rays hit the
// sphere, so 4 of 5 warps carry one or two such lanes and every one of them walks the whole sphere code for them.
// Regrouping all paths every bounce (sail_wavefront.cuh) fixes that but pays barriers and a full state exchange for
// the 95 % of paths that were coherent already.  Here only the minority moves:
//   * a CTA is SAIL_SRV_PATH_WARPS *path warps* (one thread = one pixel, as in sail_mega.cuh) plus ONE server warp;
//   * after the nearest-hit scan a path warp parks the lanes whose winner is not a cube / Cornell box: ray, distance,
//     RNG pair and pixel go into the lane's own shared-memory slot (12 words), the slot index into a CTA-wide list;
//     it signals `bar.arrive` (named barrier 1, does not wait) and shades its box lanes -- the bulk of the work;
//   * the server warp waits on barrier 1, takes up to 32 parked hits of ALL path warps per pass -- one walk through
//     the sphere / quadric code with several lanes instead of one walk per path warp with one or two --, writes
//     emission+direct, BSDF weight and the next ray back into the slots and signals barrier 2;
//   * the path warps meet barrier 2 after their own shading (by then the server is usually done) and the parked
//     lanes fold the answer into their path state with the very operations the megakernel uses.
// Which lane evaluates a hit does not change one bit of it: results equal the megakernel's and the oracle's
// (tests/test_gpu_parity.py).  The reference has no counterpart -- its divergence is the GPU driver's business
// (/root/reference/src/shader/shape/shader.shape.js:28-51 is one big per-pixel loop).
#ifndef SAIL_SERVER_CUH
#define SAIL_SERVER_CUH
#include "sail_mega.cuh"

#ifndef SAIL_SRV_PATH_WARPS
#define SAIL_SRV_PATH_WARPS 7
#endif
#define SAIL_SRV_PATHS (SAIL_SRV_PATH_WARPS * 32)
#define SAIL_SRV_THREADS (SAIL_SRV_PATHS + 32)
#define SAIL_SRV_WORDS 12
// material categories a box hit / a non-box hit can carry (sail_jit.cpp knows; ahead of time: all)
#ifndef SAIL_JIT_MATS_BOX
#define SAIL_JIT_MATS_BOX 0xffffffffu
#endif
#ifndef SAIL_JIT_MATS_OFF
#define SAIL_JIT_MATS_OFF 0xffffffffu
#endif

// Named barriers 1 (requests complete) and 2 (answers complete); every thread of the CTA takes part in both, one side
// arriving without waiting.  "memory": shared-memory traffic must not move across them.
SD void srvArrive(int id) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "n"(SAIL_SRV_THREADS) : "memory"); }
SD void srvSync(int id) { asm volatile("bar.sync %0, %1;" ::"r"(id), "n"(SAIL_SRV_THREADS) : "memory"); }
SD bool srvOffloaded(int category) { return category != SAIL_CUBE && category != SAIL_CORNELLBOX; }

template <int F>
SD void sailServerBody(const SailLaunch& L) {
    __shared__ float slot[SAIL_SRV_WORDS][SAIL_SRV_PATHS];  // SoA by path thread: lane-consecutive words
    __shared__ int list[SAIL_SRV_PATHS];                    // path threads parked this bounce
    __shared__ int count;
    const SceneView sc = stageScene(L.scene);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) count = 0;
    __syncthreads();

    if (warp == SAIL_SRV_PATH_WARPS) {
        // ---------------- server warp: one pass per 32 parked hits, every bounce of every frame ----------------
        AovSink aov;
        aov.normal = L.normal; aov.position = L.position; aov.hitId = L.hitId; aov.depth = L.depth;
        for (int k = 0; k < L.nframes; k++) {
            const bool aovOn = L.writeAov && k == L.nframes - 1;
            for (int deepth = 1; deepth <= L.maxBounces; deepth++) {
                srvSync(1);
                const int n = *(volatile int*)&count;
                for (int base = 0; base < n; base += 32) {
                    const int i = base + lane;
                    if (i < n) {
                        const int src = list[i];
                        Ray ray;
                        ray.o = mk3(slot[0][src], slot[1][src], slot[2][src]);
                        ray.d = mk3(slot[3][src], slot[4][src], slot[5][src]);
                        const float t = slot[6][src];
                        const V2 u = mk2(slot[7][src], slot[8][src]);
                        const int id = __float_as_int(slot[10][src]);
                        Hit h;
                        hitRecord<F, 2>(sc, id, t, ray, h);
                        if (deepth == 1 && aovOn) {
                            aov.pix = (size_t)(unsigned)__float_as_int(slot[9][src]);
                            writeAov(aov, h.n, h.p, id, t);
                        }
                        V3 wi = splat3(0.0f), _fpdf = splat3(0.0f), contrib, o = splat3(0.0f);
                        if (SAIL_IS_LAST_BOUNCE(deepth, L.maxBounces) && lastBounceIsEmissionOnly<F>(h)) {  // see tracePath(): dead BSDF sample
                            contrib = h.emission + splat3(0.0f);
                        } else {
                            contrib = shade<F>(sc, h, u, -ray.d, wi, _fpdf, SAIL_JIT_MATS_OFF);
                            const float outdot = dot3(h.n, wi);
                            o = h.p + h.n * (outdot > SAIL_EPSILON ? 0.0001f : -0.0001f);
                        }
                        slot[0][src] = contrib.x; slot[1][src] = contrib.y; slot[2][src] = contrib.z;
                        slot[3][src] = _fpdf.x; slot[4][src] = _fpdf.y; slot[5][src] = _fpdf.z;
                        slot[6][src] = o.x; slot[7][src] = o.y; slot[8][src] = o.z;
                        slot[9][src] = wi.x; slot[10][src] = wi.y; slot[11][src] = wi.z;
                    }
                }
                __syncwarp();
                if (lane == 0) count = 0;  // the next bounce's requests start only behind barrier 2
                asm volatile("fence.acq_rel.cta;" ::: "memory");
                srvArrive(2);
            }
        }
        return;
    }

    // ---------------- path warps: a strip of SAIL_SRV_PATH_WARPS 8x4 sub-tiles ----------------
    const int x = ((int)blockIdx.x * SAIL_SRV_PATH_WARPS + warp) * 8 + (lane & 7);
    const int r = (int)blockIdx.y * 4;  // row within this launch's rows
    int y = L.y0 + r;
    if (L.bandStride > 1) y = (L.bandFirst + (r / L.bandRows) * L.bandStride) * L.bandRows + r % L.bandRows;
    y += lane >> 3;
    const bool inside = x < L.W && y < L.y1;  // no early exit: every warp takes part in the barriers
    const size_t pix = inside ? (size_t)y * L.W + x : 0;
    const float fx = (float)x + 0.5f, fy = (float)y + 0.5f, fz = 0.5f;
    const float a = fx * L.rcpW, b = fy * L.rcpH;

    V3 c = inside ? mk3(L.accum[3 * pix], L.accum[3 * pix + 1], L.accum[3 * pix + 2]) : splat3(0.0f);
    AovSink aov;
    aov.normal = L.normal; aov.position = L.position; aov.hitId = L.hitId; aov.depth = L.depth;
    aov.pix = pix;
    for (int k = 0; k < L.nframes; k++) {
        const SailFrameDev& f = L.frames[k];
        Ray ray;
        ray.o = ld3(f.eye);
        ray.d = primaryDir(f, a, b);
        aov.on = L.writeAov && k == L.nframes - 1;
        V3 fpdf = splat3(1.0f), e = splat3(0.0f);
        bool alive = inside;
        if (inside && aov.on && L.maxBounces < 1) writeAov(aov, splat3(0.0f), splat3(0.0f), -1, SAIL_MAX_DISTANCE);
        for (int deepth = 1; deepth <= L.maxBounces; deepth++) {
            Nearest nr;
            nr.t = SAIL_MAX_DISTANCE;
            nr.id = -1;
            if (alive) {
                nr = nearestObject<F>(sc, ray);
                if (nr.id < 0) {  // path.glsl:22 `break`
                    if (deepth == 1 && aov.on) writeAov(aov, splat3(0.0f), splat3(0.0f), nr.id, nr.t);
                    alive = false;
                }
                // last bounce on an object that neither emits nor gets NEE: adds +0 to e (see tracePath())
                if (alive && SAIL_IS_LAST_BOUNCE(deepth, L.maxBounces) && !(deepth == 1 && aov.on) && !lastBounceNeedsRecord<F>(sc.shape[nr.id])) alive = false;
            }
            const float seed = f.timeSinceStart + (float)deepth;
            const bool lastNoNee = SAIL_IS_LAST_BOUNCE(deepth, L.maxBounces) && !(F & SAIL_F_NEE);  // warp-uniform: nobody needs the RNG pair
            const V2 u = lastNoNee ? mk2(0.0f, 0.0f) : random2<F>(fx, fy, fz, seed);
            const bool off = alive && srvOffloaded(sc.shape[alive ? nr.id : 0].category);
            const unsigned m = __ballot_sync(0xffffffffu, off);
            if (m) {
                const int leader = __ffs(m) - 1;
                int base = 0;
                if (lane == leader) base = atomicAdd(&count, __popc(m));
                base = __shfl_sync(0xffffffffu, base, leader);
                if (off) {
                    list[base + __popc(m & ((1u << lane) - 1u))] = tid;
                    slot[0][tid] = ray.o.x; slot[1][tid] = ray.o.y; slot[2][tid] = ray.o.z;
                    slot[3][tid] = ray.d.x; slot[4][tid] = ray.d.y; slot[5][tid] = ray.d.z;
                    slot[6][tid] = nr.t;
                    slot[7][tid] = u.x; slot[8][tid] = u.y;
                    slot[9][tid] = __int_as_float((int)(unsigned)pix);
                    slot[10][tid] = __int_as_float(nr.id);
                }
                asm volatile("fence.acq_rel.cta;" ::: "memory");
            }
            __syncwarp();
            srvArrive(1);
            if (alive && !off) {  // the coherent majority: box hit record + shading, while the server works
                Hit h;
                hitRecord<F, 1>(sc, nr.id, nr.t, ray, h);
                if (deepth == 1 && aov.on) writeAov(aov, h.n, h.p, h.index, h.d);
                if (SAIL_IS_LAST_BOUNCE(deepth, L.maxBounces) && lastBounceIsEmissionOnly<F>(h)) {
                    e = e + (h.emission + splat3(0.0f)) * fpdf;
                } else {
                    V3 wi, _fpdf;
                    e = e + shade<F>(sc, h, u, -ray.d, wi, _fpdf, SAIL_JIT_MATS_BOX) * fpdf;
                    fpdf = fpdf * _fpdf;
                    const float outdot = dot3(h.n, wi);
                    ray.o = h.p + h.n * (outdot > SAIL_EPSILON ? 0.0001f : -0.0001f);
                    ray.d = wi;
                }
            }
            __syncwarp();
            srvSync(2);
            if (off) {  // path.glsl:25-36 on the server's answer
                const V3 contrib = mk3(slot[0][tid], slot[1][tid], slot[2][tid]);
                const V3 _fpdf = mk3(slot[3][tid], slot[4][tid], slot[5][tid]);
                e = e + contrib * fpdf;
                fpdf = fpdf * _fpdf;
                ray.o = mk3(slot[6][tid], slot[7][tid], slot[8][tid]);
                ray.d = mk3(slot[9][tid], slot[10][tid], slot[11][tid]);
            }
        }
        e = mk3(nanToZero(e.x), nanToZero(e.y), nanToZero(e.z));
        if (L.accumMode == 0) c = mix3(e, c, f.textureWeight);  // fstrace.glsl:14
        else c = c + e;
    }
    if (!inside) return;
    L.accum[3 * pix] = c.x;
    L.accum[3 * pix + 1] = c.y;
    L.accum[3 * pix + 2] = c.z;
}

#endif
