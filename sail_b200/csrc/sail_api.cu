// sail_api.cu -- libsail_b200.so: the C ABI of include/sail_b200.h over the sm_100a kernels.
//
// Stands in for the reference's render orchestration + GPU abstraction on the hot path:
// Renderer/Tracer (/root/reference/src/core/renderer.js:8-73, src/core/tracer.js:8-101) and ShaderProgram's
// framebuffer/texture/uniform plumbing (src/core/webgl.js:4-135).  Device-resident FP32 buffers replace
// ShaderProgram.frameCache[0..3]; a packed scene block replaces the three R32F textures; CUDA kernels
// replace the trace and render programs.  No CPU fallback exists in this file or anywhere in the library.
//
// A context is one or several *members* (one GPU each).  The reference's single global GL context
// (src/core/webgl.js:215-227) corresponds to a one-member context.  A sharded context owns a stream,
// the full-image buffers and the communicator of every member; it splits each sail_render_n() over them by sample
// index or by row band and combines the members' buffers on the root when an image is asked for (sail_combine.cuh).
// Members are either all in this process (sail_create_info.ndev > 1) or one per process (sail_group_join).
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <new>
#include <vector>

#include "../../include/sail_b200.h"
#include "sail_combine.cuh"
#include "sail_comm.h"
#include "sail_diag.cuh"
#include "sail_jit.h"
#include "sail_mega.cuh"
#include "sail_present.cuh"
#include "sail_queue.cuh"

static_assert(SAIL_COMBINE_MAX == SAIL_MAX_DEVICES, "sail_combine.cuh and sail_b200.h disagree on the group size");

struct SailMember {
    int device = 0;
    int index = 0;  // position in the group: decides its share of the frames / row bands
    cudaStream_t stream = nullptr;
    cudaEvent_t evStart = nullptr, evStop = nullptr;
    cudaEvent_t evDone = nullptr, evCombined = nullptr;  // sums final / share delivered to the root (peer-memory combine)
    SailSceneBlock* dScene = nullptr;
    float *accum = nullptr, *normal = nullptr, *position = nullptr, *depth = nullptr;
    int* hitId = nullptr;
    bool timed = false;
    // the scene-specialised kernels loaded on this device: [kind 0 mega, 1 wavefront][variant: bit 0 = RNG sine for |arg| < 2^22 only, bit 1 = with Russian roulette]
    void* jitFn[3][4] = {};
    unsigned jitBlock[3][4] = {};
    void* jitStage[4][5] = {};  // kind 2 (queue wavefront): the five stage kernels per variant
    // queue wavefront (sail_queue.cuh): path state, permutation, counters for one batch; allocated at first use
    float4* qRay = nullptr;
    float4* qCarry = nullptr;
    int2* qHit = nullptr;
    int* qPerm = nullptr;
    int* qCounters = nullptr;
    size_t qCapacity = 0;  // paths
    sailNcclComm_t comm = nullptr;
};

struct sail_ctx {
    int W = 0, H = 0, y0 = 0, y1 = 0;
    int bandRows = 0, bandStride = 1, bandFirst = 0;  // host-managed bands of a one-member context (sail_set_row_bands)
    int nlocal = 1;                                   // members driven by this process
    SailMember mem[SAIL_MAX_DEVICES];
    // group = all members of the sharded image, here or in other processes
    int groupSize = 1, firstMember = 0;
    int shardMode = SAIL_SHARD_NONE, groupBandRows = 8, combine = SAIL_COMBINE_AUTO;
    bool joined = false;           // members live in other processes too (sail_group_join)
    uint64_t sampleIndex = 0;      // sample sharding: frames since the last restart (= samples in the sums)
    bool combinedValid = false;    // root's `combined` holds the image of the members' current buffers
    float* combined = nullptr;     // root only: the combined accumulation buffer of a sharded context
    cudaEvent_t evConsumed = nullptr, evCombStart = nullptr, evCombStop = nullptr;
    bool consumedOnce = false, combineTimed = false;
    sail_frame lastFrame;          // the AOVs of a sharded context are computed on the root from the last frame
    bool haveLastFrame = false, aovValid = false;
    char combineInfo[96];
    // double-buffered display (sail_present_begin / _end): copy stream, filter-done and copy-done events (root)
    cudaStream_t copyStream = nullptr;
    cudaEvent_t evPresented = nullptr, evCopied = nullptr;
    bool copyInFlight = false, copiedOnce = false;
    float* presentF32 = nullptr;
    unsigned char* presentU8 = nullptr;
    float4* presentPad = nullptr;   // colour map padded to 16 B per texel for the window filters and the wavelet (sail_pad_rgb_kernel), W*H float4
    float4* presentPad2 = nullptr;  // position map likewise (wavelet)
    sail_config cfg;
    bool sceneLoaded = false;
    std::vector<float> texparams, lights;  // kept for sail_upload_objects (tracer.js:25-40 re-uses them)
    int tn = 0, ln = 0;
    SailSceneBlock hostScene;
    uint64_t launches = 0;
    // scene-specialised kernels (sail_jit.cpp): requested per kernel kind after an upload, compiled on a worker thread
    SailJitJob* jitJob[3][4];
    bool jitRequested[3][4], jitFailed[3][4];
    char jitNote[3][4][320];
    char kernelInfo[400];
    char err[512];
    sail_ctx() {
        err[0] = 0; kernelInfo[0] = 0; memset(jitNote, 0, sizeof(jitNote)); combineInfo[0] = 0;
        memset(jitJob, 0, sizeof(jitJob)); memset(jitRequested, 0, sizeof(jitRequested)); memset(jitFailed, 0, sizeof(jitFailed));
        memset(&lastFrame, 0, sizeof(lastFrame));
    }
    bool grouped() const { return groupSize > 1; }
    bool isRoot() const { return firstMember == 0; }
    SailMember& root() { return mem[0]; }
};

extern "C" {
cudaError_t sail_launch_mega_g0(int F, const SailLaunch* L, dim3 grid, cudaStream_t stream);
cudaError_t sail_launch_mega_g1(int F, const SailLaunch* L, dim3 grid, cudaStream_t stream);
cudaError_t sail_launch_mega_g2(int F, const SailLaunch* L, dim3 grid, cudaStream_t stream);
cudaError_t sail_launch_mega_g3(int F, const SailLaunch* L, dim3 grid, cudaStream_t stream);
cudaError_t sail_launch_wave_g0(int F, const SailLaunch* L, dim3 grid, cudaStream_t stream);
cudaError_t sail_launch_wave_g1(int F, const SailLaunch* L, dim3 grid, cudaStream_t stream);
cudaError_t sail_launch_wave_g2(int F, const SailLaunch* L, dim3 grid, cudaStream_t stream);
cudaError_t sail_launch_wave_g3(int F, const SailLaunch* L, dim3 grid, cudaStream_t stream);
}

namespace {

cudaError_t launchMega(int F, const SailLaunch& L, dim3 grid, cudaStream_t stream) {
    switch (F >> 2) {
        case 0: return sail_launch_mega_g0(F, &L, grid, stream);
        case 1: return sail_launch_mega_g1(F, &L, grid, stream);
        case 2: return sail_launch_mega_g2(F, &L, grid, stream);
        case 3: return sail_launch_mega_g3(F, &L, grid, stream);
    }
    return cudaErrorInvalidValue;
}

cudaError_t launchWave(int F, const SailLaunch& L, dim3 grid, cudaStream_t stream) {
    switch (F >> 2) {
        case 0: return sail_launch_wave_g0(F, &L, grid, stream);
        case 1: return sail_launch_wave_g1(F, &L, grid, stream);
        case 2: return sail_launch_wave_g2(F, &L, grid, stream);
        case 3: return sail_launch_wave_g3(F, &L, grid, stream);
    }
    return cudaErrorInvalidValue;
}

int setErr(sail_ctx* c, int code, const char* fmt, ...) {
    if (c) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(c->err, sizeof(c->err), fmt, ap);
        va_end(ap);
    }
    return code;
}
#define CU(call)                                                                                          \
    do {                                                                                                  \
        cudaError_t e_ = (call);                                                                          \
        if (e_ != cudaSuccess) return setErr(ctx, SAIL_ERR_CUDA, "%s failed: %s", #call, cudaGetErrorString(e_)); \
    } while (0)
#define NC(call)                                                                                                          \
    do {                                                                                                                  \
        int r_ = (call);                                                                                                  \
        if (r_ != 0) return setErr(ctx, SAIL_ERR_CUDA, "%s failed: %s", #call, sail_nccl()->getErrorString(r_)); \
    } while (0)

size_t npix(const sail_ctx* c) { return (size_t)c->W * c->H; }

// vstrace.glsl:4-7 for the four quad corners + the edge differences of the interpolation, in the FP32
// arithmetic of the shader (this file is compiled with -ffp-contract=off on the host side).
struct H3 { float x, y, z; };
H3 cornerRay(const float* M, const float* eye, float vx, float vy) {
    const float vz = 0.0f, vw = 1.0f;
    float px = M[0] * vx + M[4] * vy + M[8] * vz + M[12] * vw;
    float py = M[1] * vx + M[5] * vy + M[9] * vz + M[13] * vw;
    float pz = M[2] * vx + M[6] * vy + M[10] * vz + M[14] * vw;
    float pw = M[3] * vx + M[7] * vy + M[11] * vz + M[15] * vw;
    float rw = dm_rcp(pw);
    H3 d{px * rw - eye[0], py * rw - eye[1], pz * rw - eye[2]};  // ensure3byW(...) - eye
    float inv = dm_rcp(dm_sqrt(d.x * d.x + d.y * d.y + d.z * d.z));
    return H3{d.x * inv, d.y * inv, d.z * inv};
}
void sub3(float* out, H3 a, H3 b) { out[0] = a.x - b.x; out[1] = a.y - b.y; out[2] = a.z - b.z; }
void lowerFrame(const sail_frame& f, SailFrameDev& o) {
    H3 d0 = cornerRay(f.matrix, f.eye, -1.0f, -1.0f), d1 = cornerRay(f.matrix, f.eye, -1.0f, 1.0f);
    H3 d2 = cornerRay(f.matrix, f.eye, 1.0f, -1.0f), d3 = cornerRay(f.matrix, f.eye, 1.0f, 1.0f);
    o.d0[0] = d0.x; o.d0[1] = d0.y; o.d0[2] = d0.z;
    o.d3[0] = d3.x; o.d3[1] = d3.y; o.d3[2] = d3.z;
    sub3(o.e20, d2, d0); sub3(o.e10, d1, d0); sub3(o.e13, d1, d3); sub3(o.e23, d2, d3);
    for (int k = 0; k < 3; k++) o.eye[k] = f.eye[k];
    o.textureWeight = f.textureWeight;
    o.timeSinceStart = f.timeSinceStart;
    o.pad = 0.0f;
}

// Commits a packed scene: only called with a block that sail_pack_scene() accepted, so a failed upload leaves the
// context (host block, device blocks, cached kernels) exactly as it was.
int uploadBlock(sail_ctx* ctx, const SailSceneBlock& blk);
int ensureAovs(sail_ctx* ctx);
int uploadBlock(sail_ctx* ctx, const SailSceneBlock& blk) {
    int arc = ensureAovs(ctx);  // the AOVs of the frames rendered so far belong to the scene that is about to be replaced
    if (arc) return arc;
    ctx->hostScene = blk;
    for (int i = 0; i < ctx->nlocal; i++) {
        SailMember& m = ctx->mem[i];
        CU(cudaSetDevice(m.device));
        CU(cudaMemcpyAsync(m.dScene, &ctx->hostScene, sizeof(SailSceneBlock), cudaMemcpyHostToDevice, m.stream));
    }
    for (int i = 0; i < ctx->nlocal; i++) {
        CU(cudaSetDevice(ctx->mem[i].device));
        CU(cudaStreamSynchronize(ctx->mem[i].stream));  // hostScene may be rewritten by the next upload
        memset(ctx->mem[i].jitFn, 0, sizeof(ctx->mem[i].jitFn));
    }
    ctx->sceneLoaded = true;
    // the structure may have changed: look the specialised kernels up again
    memset(ctx->jitRequested, 0, sizeof(ctx->jitRequested));
    memset(ctx->jitFailed, 0, sizeof(ctx->jitFailed));
    memset(ctx->jitJob, 0, sizeof(ctx->jitJob));
    ctx->aovValid = false;
    return SAIL_OK;
}

int initMember(sail_ctx* ctx, SailMember& m, int device, int index) {
    m.device = device;
    m.index = index;
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10)
        return setErr(ctx, SAIL_ERR_CUDA, "device %d is sm_%d%d; this library carries sm_100a code only", device, prop.major, prop.minor);
    CU(cudaSetDevice(device));
    CU(cudaStreamCreateWithFlags(&m.stream, cudaStreamNonBlocking));
    CU(cudaEventCreate(&m.evStart));
    CU(cudaEventCreate(&m.evStop));
    CU(cudaEventCreateWithFlags(&m.evDone, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&m.evCombined, cudaEventDisableTiming));
    size_t n = npix(ctx);
    CU(cudaMalloc(&m.dScene, sizeof(SailSceneBlock)));
    CU(cudaMalloc(&m.accum, n * 3 * sizeof(float)));
    CU(cudaMalloc(&m.normal, n * 3 * sizeof(float)));
    CU(cudaMalloc(&m.position, n * 3 * sizeof(float)));
    CU(cudaMalloc(&m.depth, n * sizeof(float)));
    CU(cudaMalloc(&m.hitId, n * sizeof(int)));
    return SAIL_OK;
}

// Buffers and events that exist once per context, on the root's device.
int initRoot(sail_ctx* ctx) {
    if (!ctx->isRoot()) return SAIL_OK;
    SailMember& r = ctx->root();
    CU(cudaSetDevice(r.device));
    if (!ctx->copyStream) {
        CU(cudaStreamCreateWithFlags(&ctx->copyStream, cudaStreamNonBlocking));
        CU(cudaEventCreateWithFlags(&ctx->evPresented, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&ctx->evCopied, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&ctx->evConsumed, cudaEventDisableTiming));
        CU(cudaEventCreate(&ctx->evCombStart));
        CU(cudaEventCreate(&ctx->evCombStop));
    }
    if (ctx->grouped() && !ctx->combined) CU(cudaMalloc(&ctx->combined, npix(ctx) * 3 * sizeof(float)));
    return SAIL_OK;
}

int checkShard(sail_ctx* ctx, int shard_mode, int band_rows, int combine) {
    if (shard_mode != SAIL_SHARD_SAMPLE && shard_mode != SAIL_SHARD_TILE)
        return setErr(ctx, SAIL_ERR_INVALID, "a multi-GPU context needs shard_mode SAIL_SHARD_SAMPLE or SAIL_SHARD_TILE (got %d)", shard_mode);
    if (band_rows < 0 || band_rows % SAIL_TILE_H) return setErr(ctx, SAIL_ERR_INVALID, "band_rows %d must be a multiple of %d", band_rows, SAIL_TILE_H);
    if (combine < SAIL_COMBINE_AUTO || combine > SAIL_COMBINE_P2P) return setErr(ctx, SAIL_ERR_INVALID, "unknown combine mode %d", combine);
    ctx->shardMode = shard_mode;
    ctx->groupBandRows = band_rows ? band_rows : 8;  // 8 / 16 / 32 rows on 8 GPUs, config 5: 10.2 / 9.6 / 8.6 Gpaths/s (slowest rank +7 / +14 / +28 %)
    return SAIL_OK;
}

// Peer mappings between all local members (both directions).  Returns false when some pair cannot map each other.
bool enablePeerAccess(sail_ctx* ctx) {
    for (int a = 0; a < ctx->nlocal; a++)
        for (int b = 0; b < ctx->nlocal; b++) {
            if (a == b || ctx->mem[a].device == ctx->mem[b].device) continue;  // members sharing a GPU see each other's memory anyway
            int can = 0;
            if (cudaDeviceCanAccessPeer(&can, ctx->mem[a].device, ctx->mem[b].device) != cudaSuccess || !can) { cudaGetLastError(); return false; }
        }
    for (int a = 0; a < ctx->nlocal; a++) {
        if (cudaSetDevice(ctx->mem[a].device) != cudaSuccess) return false;
        for (int b = 0; b < ctx->nlocal; b++) {
            if (a == b || ctx->mem[a].device == ctx->mem[b].device) continue;
            cudaError_t e = cudaDeviceEnablePeerAccess(ctx->mem[b].device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); return false; }
            cudaGetLastError();
        }
    }
    return true;
}

int resetAll(sail_ctx* ctx) {
    size_t n = npix(ctx);
    for (int i = 0; i < ctx->nlocal; i++) {
        SailMember& m = ctx->mem[i];
        CU(cudaSetDevice(m.device));
        CU(cudaMemsetAsync(m.accum, 0, n * 3 * sizeof(float), m.stream));
        CU(cudaMemsetAsync(m.normal, 0, n * 3 * sizeof(float), m.stream));
        CU(cudaMemsetAsync(m.position, 0, n * 3 * sizeof(float), m.stream));
        CU(cudaMemsetAsync(m.depth, 0, n * sizeof(float), m.stream));
        CU(cudaMemsetAsync(m.hitId, 0xff, n * sizeof(int), m.stream));
    }
    ctx->sampleIndex = 0;
    ctx->combinedValid = false;
    ctx->aovValid = false;
    ctx->haveLastFrame = false;
    return SAIL_OK;
}

// Rows of tile-sharding band b, and how many bands member g owns.
int bandCount(const sail_ctx* ctx) { return (ctx->H + ctx->groupBandRows - 1) / ctx->groupBandRows; }
int bandsOf(const sail_ctx* ctx, int g) {
    const int nb = bandCount(ctx);
    return g < nb ? (nb - g + ctx->groupSize - 1) / ctx->groupSize : 0;
}

// The peer-memory combine of a context whose members all live in this process (sail_combine.cuh).
int combineP2P(sail_ctx* ctx) {
    SailMember& root = ctx->root();
    const size_t count = npix(ctx) * 3;
    const int N = ctx->nlocal;
    for (int i = 0; i < N; i++) {
        CU(cudaSetDevice(ctx->mem[i].device));
        CU(cudaEventRecord(ctx->mem[i].evDone, ctx->mem[i].stream));
    }
    for (int g = 0; g < N; g++) {
        SailMember& m = ctx->mem[g];
        CU(cudaSetDevice(m.device));
        // the previous image must have left the root's combined buffer before anyone writes the next one into it
        if (g != 0 && ctx->consumedOnce) CU(cudaStreamWaitEvent(m.stream, ctx->evConsumed, 0));
        if (ctx->shardMode == SAIL_SHARD_SAMPLE) {
            for (int o = 0; o < N; o++)
                if (o != g) CU(cudaStreamWaitEvent(m.stream, ctx->mem[o].evDone, 0));
            SailCombineParams P;
            for (int o = 0; o < SAIL_COMBINE_MAX; o++) P.src[o] = o < N ? ctx->mem[o].accum : nullptr;
            P.dst = ctx->combined;
            P.nsrc = N;
            const size_t per = ((count / 4 + N - 1) / N) * 4;  // floats per slice, a multiple of 4
            P.begin = (unsigned long long)(per * g < count ? per * g : count);
            P.end = (unsigned long long)((g == N - 1 || per * (g + 1) > count) ? count : per * (g + 1));
            P.scale = (ctx->cfg.accum_mode == SAIL_ACCUM_MEAN && ctx->sampleIndex > 0) ? dm_rcp((float)ctx->sampleIndex) : 1.0f;
            if (P.end > P.begin) {
                const size_t nvec = (P.end - P.begin) / 4;
                unsigned blocks = (unsigned)((nvec + 255) / 256);
                if (blocks > 148u * 8u) blocks = 148u * 8u;
                if (blocks < 1) blocks = 1;
                sail_combine_sum_kernel<<<blocks, 256, 0, m.stream>>>(P);
                ctx->launches++;
                CU(cudaGetLastError());
            }
        } else {  // tile sharding: this member's row bands, one strided copy (+ a short last band)
            const size_t rowBytes = (size_t)ctx->W * 3 * sizeof(float), bandBytes = rowBytes * ctx->groupBandRows;
            const int nb = bandCount(ctx), mine = bandsOf(ctx, g);
            if (mine > 0) {
                const int lastBand = g + (mine - 1) * N;
                const bool lastShort = lastBand == nb - 1 && (ctx->H % ctx->groupBandRows) != 0;
                const int full = lastShort ? mine - 1 : mine;
                const size_t off = (size_t)g * bandBytes;
                if (full > 0)
                    CU(cudaMemcpy2DAsync((char*)ctx->combined + off, bandBytes * N, (const char*)m.accum + off, bandBytes * N, bandBytes, (size_t)full,
                                         cudaMemcpyDefault, m.stream));
                if (lastShort) {
                    const size_t o2 = (size_t)lastBand * bandBytes;
                    CU(cudaMemcpyAsync((char*)ctx->combined + o2, (const char*)m.accum + o2, rowBytes * (size_t)(ctx->H % ctx->groupBandRows), cudaMemcpyDefault, m.stream));
                }
            }
        }
        CU(cudaEventRecord(m.evCombined, m.stream));
    }
    // the root consumes the image after every share has arrived; in sample mode a member may not start changing its
    // sums again before all peers have read them
    for (int g = 0; g < N; g++) {
        if (g != 0 && ctx->shardMode != SAIL_SHARD_SAMPLE) continue;
        CU(cudaSetDevice(ctx->mem[g].device));
        for (int o = 0; o < N; o++)
            if (o != g) CU(cudaStreamWaitEvent(ctx->mem[g].stream, ctx->mem[o].evCombined, 0));
    }
    (void)root;
    return SAIL_OK;
}

// ncclReduce(sum) of every member's buffer to the root; tile sharding uses it too (rows a member does not own are zero).
int combineNccl(sail_ctx* ctx) {
    const SailNccl* nc = sail_nccl();
    const size_t count = npix(ctx) * 3;
    NC(nc->groupStart());
    for (int i = 0; i < ctx->nlocal; i++) {
        SailMember& m = ctx->mem[i];
        float* recv = (ctx->isRoot() && i == 0) ? ctx->combined : m.accum;  // only read on the root
        int r = nc->reduce(m.accum, recv, count, SAIL_NCCL_FLOAT32, SAIL_NCCL_SUM, 0, m.comm, m.stream);
        if (r != 0) { nc->groupEnd(); return setErr(ctx, SAIL_ERR_CUDA, "ncclReduce failed: %s", nc->getErrorString(r)); }
    }
    NC(nc->groupEnd());
    if (ctx->isRoot() && ctx->shardMode == SAIL_SHARD_SAMPLE && ctx->cfg.accum_mode == SAIL_ACCUM_MEAN && ctx->sampleIndex > 0) {
        CU(cudaSetDevice(ctx->root().device));
        sail_scale_kernel<<<(unsigned)((count + 255) / 256), 256, 0, ctx->root().stream>>>(ctx->combined, count, dm_rcp((float)ctx->sampleIndex));
        ctx->launches++;
        CU(cudaGetLastError());
    }
    return SAIL_OK;
}

// Makes the root's image buffer current.  Collective on a joined context.  Returns the buffer through *img (null off-root).
int ensureCombined(sail_ctx* ctx, const float** img) {
    if (!ctx->grouped()) { *img = ctx->root().accum; return SAIL_OK; }
    *img = ctx->isRoot() ? ctx->combined : nullptr;
    if (ctx->combinedValid) return SAIL_OK;
    if (ctx->isRoot()) {
        CU(cudaSetDevice(ctx->root().device));
        CU(cudaEventRecord(ctx->evCombStart, ctx->root().stream));
    }
    int rc = ctx->combine == SAIL_COMBINE_P2P ? combineP2P(ctx) : combineNccl(ctx);
    if (rc) return rc;
    if (ctx->isRoot()) {
        CU(cudaSetDevice(ctx->root().device));
        CU(cudaEventRecord(ctx->evCombStop, ctx->root().stream));
        ctx->combineTimed = true;
    }
    ctx->combinedValid = true;
    return SAIL_OK;
}

// The first-hit AOVs of the last frame (fstrace.glsl:15-16), computed when somebody asks for them -- a reader, a filter that
// samples them, or an upload that is about to replace the scene they belong to -- instead of on every launch (32 B per pixel).
// Sharded contexts: on the root, whichever member traced that frame.
int ensureAovs(sail_ctx* ctx) {
    if (ctx->aovValid || !ctx->isRoot() || !ctx->haveLastFrame || !ctx->sceneLoaded) return SAIL_OK;
    SailMember& r = ctx->root();
    CU(cudaSetDevice(r.device));
    SailAovLaunch L;
    L.W = ctx->W; L.H = ctx->H; L.maxBounces = ctx->cfg.max_bounces;
    L.rcpW = dm_rcp((float)ctx->W); L.rcpH = dm_rcp((float)ctx->H);
    L.scene = r.dScene;
    L.normal = r.normal; L.position = r.position; L.hitId = r.hitId; L.depth = r.depth;
    lowerFrame(ctx->lastFrame, L.frame);
    dim3 grid((ctx->W + SAIL_TILE_W - 1) / SAIL_TILE_W, (ctx->H + SAIL_TILE_H - 1) / SAIL_TILE_H);
    sail_aov_kernel<<<grid, SAIL_CTA_THREADS, sailSceneSmemBytes(ctx->hostScene.n, ctx->hostScene.ln), r.stream>>>(L);
    ctx->launches++;
    CU(cudaGetLastError());
    ctx->aovValid = true;
    return SAIL_OK;
}

bool filterNeedsAovs(int id) { return id == SAIL_FILTER_NORMAL || id == SAIL_FILTER_POSITION || id == SAIL_FILTER_WAVELET; }

// The root stream has consumed the combined buffer up to here (filter or read-back queued).
int markConsumed(sail_ctx* ctx) {
    if (ctx->grouped() && ctx->isRoot()) {
        CU(cudaEventRecord(ctx->evConsumed, ctx->root().stream));
        ctx->consumedOnce = true;
    }
    return SAIL_OK;
}

void fillPresentParams(sail_ctx* ctx, PresentParams& P, const float* color) {
    P.W = ctx->W; P.H = ctx->H;
    P.filter_id = ctx->cfg.filter_id;
    P.gamma_c = ctx->cfg.filter_gamma_c;
    P.radius[0] = ctx->cfg.filter_radius[0]; P.radius[1] = ctx->cfg.filter_radius[1];
    memcpy(P.weights, ctx->cfg.filter_weights, sizeof(P.weights));
    P.color = color; P.normal = ctx->root().normal; P.position = ctx->root().position;
    P.out_f32 = nullptr;
    P.out_u8 = nullptr;
}

// The kernel kind (0 mega, 1 wavefront) SAIL_KERNEL_AUTO picks, from the measured matrix: the wavefront
// (512 paths per CTA when compiled at run time) pays where the scene has many shape codes (quadrics); among the
// ahead-of-time variants also where it has many BSDFs (metal / glass).
int kernelKind(const sail_ctx* ctx, int F, bool jit) {
    if (ctx->cfg.kernel == SAIL_KERNEL_QUEUE) {
        // whole images only (its batches are runs of consecutive pixels), scene-specialised only; else the shared-memory wavefront
        const bool whole = ctx->y0 == 0 && ctx->y1 == ctx->H && ctx->bandStride <= 1 && !(ctx->grouped() && ctx->shardMode == SAIL_SHARD_TILE);
        return (jit && whole && ctx->cfg.max_bounces <= SAIL_Q_MAX_BOUNCES && ctx->cfg.max_bounces >= 1) ? 2 : 1;
    }
    if (ctx->cfg.kernel == SAIL_KERNEL_WAVEFRONT) return 1;
    if (ctx->cfg.kernel == SAIL_KERNEL_MEGA) return 0;
    if (jit) return (F & SAIL_F_QUADRICS) ? 1 : 0;
    return (F & (SAIL_F_QUADRICS | SAIL_F_GLOSSY)) ? 1 : 0;
}

// Looks after the scene-specialised kernel of `kind`: requests its compilation once per upload and, when the cubin is
// there, loads it on every local member.  Returns true when all members have it.  `wait` blocks on the compiler.
// `small`: the kernel variant, bit 0 = RNG arguments proven < 2^22, bit 1 = Russian roulette compiled in
bool specialisedReady(sail_ctx* ctx, int kind, int small, int F, bool wait) {
    if (ctx->jitFailed[kind][small]) return false;
    // SAIL_B200_JIT_SYNC=1: SAIL_SPECIALISE_AUTO waits for the compiler too (test suites that must exercise the specialised kernels)
    const char* sync = getenv("SAIL_B200_JIT_SYNC");
    const bool alwaysWait = sync && sync[0] == '1';
    wait = wait || alwaysWait;
    char* note = ctx->jitNote[kind][small];
    const size_t noteLen = sizeof(ctx->jitNote[kind][small]);
    if (!ctx->jitRequested[kind][small]) {
        ctx->jitRequested[kind][small] = true;
        char why[300];
        why[0] = 0;
        ctx->jitJob[kind][small] = sail_jit_request(kind, F, SAIL_MIN_CTAS(F), small, &ctx->hostScene, why, sizeof(why));
        if (!ctx->jitJob[kind][small]) {
            ctx->jitFailed[kind][small] = true;
            snprintf(note, noteLen, "jit: %s", why);
            return false;
        }
        snprintf(note, noteLen, "jit: compiling");
    }
    bool all = true;
    for (int i = 0; i < ctx->nlocal; i++) {
        SailMember& m = ctx->mem[i];
        if (m.jitFn[kind][small]) continue;
        SailJitKernel k;
        char why[300];
        why[0] = 0;
        if (cudaSetDevice(m.device) != cudaSuccess) { cudaGetLastError(); all = false; continue; }
        const int rc = sail_jit_poll(ctx->jitJob[kind][small], m.device, wait ? 1 : 0, &k, why, sizeof(why));
        if (rc == 0) {
            m.jitFn[kind][small] = k.fn;
            m.jitBlock[kind][small] = k.block;
            if (kind == 2) for (int st = 0; st < 5; st++) m.jitStage[small][st] = k.stage[st];
            snprintf(note, noteLen, "%s", k.info);
        } else if (rc < 0) {
            ctx->jitFailed[kind][small] = true;
            snprintf(note, noteLen, "jit: %s", why);
            return false;
        } else {
            all = false;
        }
    }
    return all;
}

// The hash RNG takes sin(dot(gl_FragCoord.xyz + seed, K)) with seed = timeSinceStart + bounce (src/shader/util/random.glsl:1-18,
// src/shader/trace/path.glsl:21).  True when that argument stays below 2^22 -- the range of the fast reduction in detmath.h --
// for every pixel, bounce and frame of the call: all terms are positive, so the corner pixel with the largest seed bounds it.
// (About the first 95 minutes of a session at 4K; afterwards the kernel with the full-range sine runs.)
bool rngArgsSmall(const sail_ctx* ctx, const sail_frame* frames, int count) {
    float tmax = 0.0f;
    for (int k = 0; k < count; k++) tmax = frames[k].timeSinceStart > tmax ? frames[k].timeSinceStart : tmax;
    const double seed = (double)tmax + (double)(ctx->cfg.max_bounces > 1 ? ctx->cfg.max_bounces : 1);
    const double ax = (double)ctx->W + seed, ay = (double)ctx->H + seed, az = 0.5 + seed;
    const double d1 = ax * 12.9898 + ay * 78.233 + az * 151.7182, d2 = ax * 63.7264 + ay * 10.873 + az * 623.6736;
    return (d1 > d2 ? d1 : d2) * 1.001 < 4194304.0;
}

// One sail_render_n() share of a member through the queue wavefront: batches of consecutive pixels x all frames of the call.
int launchQueue(sail_ctx* ctx, SailMember& m, int small, const SailFrameDev* frames, int nframes, int accumMode) {
    const size_t npixTotal = npix(ctx);
    size_t maxPaths = (size_t)1 << 22;
    if (const char* e = getenv("SAIL_B200_QUEUE_PATHS")) {  // batch size sweep (paths per batch)
        const long long v = atoll(e);
        if (v >= 1024 && v <= ((long long)1 << 26)) maxPaths = (size_t)v;
    }
    if (!m.qRay) {
        size_t cap = npixTotal * SAIL_MAX_BATCH;
        if (cap > maxPaths) cap = maxPaths;
        if (cap < 1024) cap = 1024;
        // all or nothing: a half-allocated set must not look complete to the next call (m.qRay is the "allocated" flag)
        cudaError_t qe = cudaMalloc(&m.qCarry, cap * 2 * sizeof(float4));
        if (qe == cudaSuccess) qe = cudaMalloc(&m.qHit, cap * sizeof(int2));
        if (qe == cudaSuccess) qe = cudaMalloc(&m.qPerm, cap * sizeof(int));
        if (qe == cudaSuccess) qe = cudaMalloc(&m.qCounters, (SAIL_Q_MAX_BOUNCES + 1) * 2 * SAIL_Q_CLASSES * sizeof(int));
        if (qe == cudaSuccess) qe = cudaMalloc(&m.qRay, cap * 2 * sizeof(float4));
        if (qe != cudaSuccess) {
            cudaFree(m.qCarry); cudaFree(m.qHit); cudaFree(m.qPerm); cudaFree(m.qCounters);
            m.qCarry = nullptr; m.qHit = nullptr; m.qPerm = nullptr; m.qCounters = nullptr; m.qRay = nullptr;
            cudaGetLastError();
            return setErr(ctx, SAIL_ERR_NOMEM, "queue wavefront: %s allocating the queues for %zu paths", cudaGetErrorString(qe), cap);
        }
        m.qCapacity = cap;
    }
    size_t chunk = m.qCapacity / (size_t)nframes;
    if (chunk >= 32) chunk &= ~(size_t)31;
    if (chunk < 1) return setErr(ctx, SAIL_ERR_NOMEM, "queue wavefront: batch too small");
    SailQLaunch Q;
    Q.W = ctx->W; Q.H = ctx->H;
    Q.nframes = nframes; Q.maxBounces = ctx->cfg.max_bounces; Q.rrStart = ctx->cfg.rr_start_bounce; Q.accumMode = accumMode;
    Q.pad0 = 0;
    Q.rcpW = dm_rcp((float)ctx->W); Q.rcpH = dm_rcp((float)ctx->H);
    Q.accum = m.accum; Q.scene = m.dScene; Q.ray = m.qRay; Q.carry = m.qCarry; Q.hit = m.qHit; Q.perm = m.qPerm; Q.counters = m.qCounters;
    memcpy(Q.frames, frames, (size_t)nframes * sizeof(SailFrameDev));
    const unsigned smem = (unsigned)sailSceneSmemBytes(ctx->hostScene.n, ctx->hostScene.ln);
    char why[200];
    for (size_t pix0 = 0; pix0 < npixTotal; pix0 += chunk) {
        Q.pix0 = (int)pix0;
        Q.npix = (int)(npixTotal - pix0 < chunk ? npixTotal - pix0 : chunk);
        const unsigned pathBlocks = (unsigned)(((size_t)Q.npix * nframes + SAIL_Q_THREADS - 1) / SAIL_Q_THREADS);
        const unsigned pixBlocks = (unsigned)((Q.npix + SAIL_Q_THREADS - 1) / SAIL_Q_THREADS);
#define QSTAGE(st, blocks, sm)                                                                                                       \
    do {                                                                                                                             \
        if (sail_jit_launch(m.jitStage[small][st], &Q, blocks, 1, SAIL_Q_THREADS, sm, m.stream, why, sizeof(why)) != 0)              \
            return setErr(ctx, SAIL_ERR_CUDA, "%s", why);                                                                            \
        ctx->launches++;                                                                                                             \
    } while (0)
        Q.bounce = 0;
        QSTAGE(0, pathBlocks, 0);
        for (int d = 1; d <= Q.maxBounces; d++) {
            Q.bounce = d;
            QSTAGE(1, pathBlocks, smem);
            if (d > 1) QSTAGE(2, pathBlocks, 0);
            QSTAGE(3, pathBlocks, smem);
        }
        QSTAGE(4, pixBlocks, 0);
#undef QSTAGE
    }
    return SAIL_OK;
}

}  // namespace

extern "C" {

const char* sail_version(void) { return "sail-b200 0.2 (sm_100a)"; }
int sail_abi_version(void) { return SAIL_ABI_VERSION; }

int sail_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    int ok = 0;
    for (int i = 0; i < n; i++) {
        cudaDeviceProp p;
        if (cudaGetDeviceProperties(&p, i) == cudaSuccess && p.major == 10) ok++;
    }
    return ok;
}

void sail_default_config(sail_config* cfg) {
    if (!cfg) return;
    memset(cfg, 0, sizeof(*cfg));
    cfg->shape_mask = cfg->material_mask = cfg->texture_mask = cfg->light_mask = 0xffffffffu;
    cfg->trace_id = 0;
    cfg->max_bounces = 5;  // const/define.glsl:6
    cfg->accum_mode = SAIL_ACCUM_MEAN;
    cfg->kernel = SAIL_KERNEL_AUTO;
    cfg->filter_id = SAIL_FILTER_COLOR;
    cfg->filter_gamma_c = 2.2f;
}

int sail_create(const sail_create_info* info, sail_ctx** out) {
    if (!info || !out) return SAIL_ERR_INVALID;
    *out = nullptr;
    sail_ctx* ctx = new (std::nothrow) sail_ctx();
    if (!ctx) return SAIL_ERR_NOMEM;
    *out = ctx;  // returned even on failure so the caller can read sail_last_error(); destroy it afterwards
    if (info->width <= 0 || info->height <= 0 || (size_t)info->width * info->height > ((size_t)1 << 28))
        return setErr(ctx, SAIL_ERR_INVALID, "bad image size %dx%d", info->width, info->height);
    ctx->W = info->width;
    ctx->H = info->height;
    ctx->y0 = info->row_begin;
    ctx->y1 = info->row_end;
    if (ctx->y0 == 0 && ctx->y1 == 0) ctx->y1 = ctx->H;
    if (ctx->y0 < 0 || ctx->y1 > ctx->H || ctx->y0 >= ctx->y1)
        return setErr(ctx, SAIL_ERR_INVALID, "bad row range [%d,%d) for height %d", ctx->y0, ctx->y1, ctx->H);
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return setErr(ctx, SAIL_ERR_CUDA, "no CUDA device (%s); libsail_b200 has no CPU path", e == cudaSuccess ? "count = 0" : cudaGetErrorString(e));
    }
    sail_default_config(&ctx->cfg);
    const int want = info->ndev <= 1 ? 1 : info->ndev;
    if (want > SAIL_MAX_DEVICES) return setErr(ctx, SAIL_ERR_INVALID, "ndev %d exceeds SAIL_MAX_DEVICES (%d)", want, SAIL_MAX_DEVICES);
    bool shared = false;
    if (want > 1) {
        if (ctx->y0 != 0 || ctx->y1 != ctx->H) return setErr(ctx, SAIL_ERR_INVALID, "a multi-GPU context shards the image itself: row_begin / row_end must be 0");
        int rc = checkShard(ctx, info->shard_mode, info->band_rows, info->combine);
        if (rc) return rc;
        for (int i = 0; i < want; i++) {
            if (info->device_ids[i] < 0 || info->device_ids[i] >= ndev)
                return setErr(ctx, SAIL_ERR_INVALID, "device_ids[%d] = %d out of range (%d devices)", i, info->device_ids[i], ndev);
            // A device may be listed more than once: several members then share that GPU, each with its own stream and buffers
            // (no speed-up -- it lets a one-GPU machine run, and test, the sharded path); they combine through peer memory, NCCL
            // refuses two ranks on one device.
            for (int j = 0; j < i; j++)
                if (info->device_ids[j] == info->device_ids[i]) shared = true;
        }
        if (shared && info->combine == SAIL_COMBINE_NCCL)
            return setErr(ctx, SAIL_ERR_INVALID, "SAIL_COMBINE_NCCL needs distinct devices (a device is listed twice); use SAIL_COMBINE_AUTO / _P2P");
    } else if (info->device < 0 || info->device >= ndev) {
        return setErr(ctx, SAIL_ERR_INVALID, "device %d out of range (%d devices)", info->device, ndev);
    }
    ctx->nlocal = 0;
    for (int i = 0; i < want; i++) {
        int rc = initMember(ctx, ctx->mem[i], want > 1 ? info->device_ids[i] : info->device, i);
        ctx->nlocal = i + 1;  // destroyed with the context even when half initialised
        if (rc) return rc;
    }
    ctx->groupSize = want;
    ctx->firstMember = 0;
    int rc = initRoot(ctx);
    if (rc) return rc;
    if (want > 1) {
        const bool peers = enablePeerAccess(ctx);
        int combine = info->combine;
        if (combine == SAIL_COMBINE_P2P && !peers) return setErr(ctx, SAIL_ERR_CUDA, "SAIL_COMBINE_P2P: the devices cannot map each other's memory");
        if (combine == SAIL_COMBINE_AUTO) combine = peers ? SAIL_COMBINE_P2P : SAIL_COMBINE_NCCL;
        if (combine == SAIL_COMBINE_NCCL && shared) return setErr(ctx, SAIL_ERR_CUDA, "members share a GPU and the other devices cannot map each other's memory: no combine path");
        if (combine == SAIL_COMBINE_NCCL) {
            const SailNccl* nc = sail_nccl();
            if (!nc->ok) return setErr(ctx, SAIL_ERR_CUDA, "%s", nc->why);
            sailNcclComm_t comms[SAIL_MAX_DEVICES];
            int devs[SAIL_MAX_DEVICES];
            for (int i = 0; i < want; i++) devs[i] = ctx->mem[i].device;
            NC(nc->commInitAll(comms, want, devs));
            for (int i = 0; i < want; i++) ctx->mem[i].comm = comms[i];
            snprintf(ctx->combineInfo, sizeof(ctx->combineInfo), "nccl %d.%d.%d reduce", nc->version / 10000, (nc->version / 100) % 100, nc->version % 100);
        } else {
            snprintf(ctx->combineInfo, sizeof(ctx->combineInfo), "p2p %s", ctx->shardMode == SAIL_SHARD_SAMPLE ? "slice-sum kernels" : "band copies");
        }
        ctx->combine = combine;
    }
    return resetAll(ctx);
}

void sail_destroy(sail_ctx* ctx) {
    if (!ctx) return;
    for (int i = 0; i < ctx->nlocal; i++) {
        SailMember& m = ctx->mem[i];
        if (!m.stream) continue;
        cudaSetDevice(m.device);
        cudaStreamSynchronize(m.stream);
        if (i == 0 && ctx->copyStream) cudaStreamSynchronize(ctx->copyStream);
    }
    for (int i = 0; i < ctx->nlocal; i++) {
        SailMember& m = ctx->mem[i];
        cudaSetDevice(m.device);
        if (m.comm && sail_nccl()->ok) sail_nccl()->commDestroy(m.comm);
        cudaFree(m.dScene); cudaFree(m.accum); cudaFree(m.normal); cudaFree(m.position); cudaFree(m.depth); cudaFree(m.hitId);
        cudaFree(m.qRay); cudaFree(m.qCarry); cudaFree(m.qHit); cudaFree(m.qPerm); cudaFree(m.qCounters);
        if (m.evStart) cudaEventDestroy(m.evStart);
        if (m.evStop) cudaEventDestroy(m.evStop);
        if (m.evDone) cudaEventDestroy(m.evDone);
        if (m.evCombined) cudaEventDestroy(m.evCombined);
        if (i == 0) {
            cudaFree(ctx->combined); cudaFree(ctx->presentF32); cudaFree(ctx->presentU8); cudaFree(ctx->presentPad); cudaFree(ctx->presentPad2);
            if (ctx->evPresented) cudaEventDestroy(ctx->evPresented);
            if (ctx->evCopied) cudaEventDestroy(ctx->evCopied);
            if (ctx->evConsumed) cudaEventDestroy(ctx->evConsumed);
            if (ctx->evCombStart) cudaEventDestroy(ctx->evCombStart);
            if (ctx->evCombStop) cudaEventDestroy(ctx->evCombStop);
            if (ctx->copyStream) cudaStreamDestroy(ctx->copyStream);
        }
        if (m.stream) cudaStreamDestroy(m.stream);
    }
    cudaGetLastError();
    delete ctx;
}

const char* sail_last_error(const sail_ctx* ctx) { return ctx ? ctx->err : "null context"; }

int sail_group_id(uint8_t id[SAIL_GROUP_ID_BYTES]) {
    if (!id) return SAIL_ERR_INVALID;
    const SailNccl* nc = sail_nccl();
    if (!nc->ok) return SAIL_ERR_CUDA;
    SailNcclId u;
    if (nc->getUniqueId(&u) != 0) return SAIL_ERR_CUDA;
    static_assert(sizeof(u) == SAIL_GROUP_ID_BYTES, "ncclUniqueId size");
    memcpy(id, &u, sizeof(u));
    return SAIL_OK;
}

int sail_group_join(sail_ctx* ctx, int32_t nranks, int32_t rank, const uint8_t id[SAIL_GROUP_ID_BYTES], int32_t shard_mode,
                    int32_t band_rows, int32_t combine) {
    if (!ctx || !id) return SAIL_ERR_INVALID;
    if (!ctx->mem[0].accum) return setErr(ctx, SAIL_ERR_STATE, "context was not created successfully");
    if (ctx->grouped()) return setErr(ctx, SAIL_ERR_STATE, "the context already is part of a group");
    if (nranks < 2 || nranks > SAIL_MAX_DEVICES || rank < 0 || rank >= nranks) return setErr(ctx, SAIL_ERR_INVALID, "bad group: rank %d of %d (2..%d ranks)", rank, nranks, SAIL_MAX_DEVICES);
    if (ctx->y0 != 0 || ctx->y1 != ctx->H || ctx->bandStride > 1) return setErr(ctx, SAIL_ERR_INVALID, "a group shards the image itself: the context must cover all rows");
    int rc = checkShard(ctx, shard_mode, band_rows, combine);
    if (rc) return rc;
    if (combine == SAIL_COMBINE_P2P) return setErr(ctx, SAIL_ERR_INVALID, "SAIL_COMBINE_P2P needs all members in one process (sail_create_info.ndev); a joined group combines with NCCL");
    const SailNccl* nc = sail_nccl();
    if (!nc->ok) return setErr(ctx, SAIL_ERR_CUDA, "%s", nc->why);
    SailNcclId u;
    memcpy(&u, id, sizeof(u));
    CU(cudaSetDevice(ctx->mem[0].device));
    NC(nc->commInitRank(&ctx->mem[0].comm, nranks, u, rank));
    ctx->groupSize = nranks;
    ctx->firstMember = rank;
    ctx->mem[0].index = rank;
    ctx->joined = true;
    ctx->combine = SAIL_COMBINE_NCCL;
    snprintf(ctx->combineInfo, sizeof(ctx->combineInfo), "nccl %d.%d.%d reduce", nc->version / 10000, (nc->version / 100) % 100, nc->version % 100);
    rc = initRoot(ctx);
    if (rc) return rc;
    return resetAll(ctx);
}

int sail_group_info(sail_ctx* ctx, int32_t* nmembers, int32_t* first_member, char* buf, int32_t len) {
    if (!ctx) return SAIL_ERR_INVALID;
    if (nmembers) *nmembers = ctx->groupSize;
    if (first_member) *first_member = ctx->firstMember;
    if (buf && len > 0) {
        if (ctx->grouped()) snprintf(buf, (size_t)len, "%s-sharded over %d GPUs (%d in this process), combine: %s", ctx->shardMode == SAIL_SHARD_SAMPLE ? "sample" : "tile", ctx->groupSize, ctx->nlocal, ctx->combineInfo);
        else snprintf(buf, (size_t)len, "single GPU");
    }
    return SAIL_OK;
}

int sail_configure(sail_ctx* ctx, const sail_config* cfg) {
    if (!ctx || !cfg) return SAIL_ERR_INVALID;
    if (cfg->trace_id != 0) return setErr(ctx, SAIL_ERR_INVALID, "unknown trace plugin id %d (only 0 = path)", cfg->trace_id);
    if (cfg->max_bounces < 0 || cfg->max_bounces > 1024) return setErr(ctx, SAIL_ERR_INVALID, "max_bounces %d out of range", cfg->max_bounces);
    if (cfg->filter_id < 0 || cfg->filter_id > SAIL_FILTER_TONEMAPPING) return setErr(ctx, SAIL_ERR_INVALID, "unknown filter id %d", cfg->filter_id);
    if (cfg->accum_mode != SAIL_ACCUM_MEAN && cfg->accum_mode != SAIL_ACCUM_SUM) return setErr(ctx, SAIL_ERR_INVALID, "unknown accum_mode %d", cfg->accum_mode);
    if (cfg->kernel < SAIL_KERNEL_MEGA || cfg->kernel > SAIL_KERNEL_QUEUE) return setErr(ctx, SAIL_ERR_INVALID, "unknown kernel %d", cfg->kernel);
    if (cfg->rr_start_bounce < 0 || cfg->rr_start_bounce > 1024) return setErr(ctx, SAIL_ERR_INVALID, "rr_start_bounce %d out of range", cfg->rr_start_bounce);
    if (cfg->specialise < SAIL_SPECIALISE_AUTO || cfg->specialise > SAIL_SPECIALISE_REQUIRE) return setErr(ctx, SAIL_ERR_INVALID, "unknown specialise mode %d", cfg->specialise);
    if (cfg->max_bounces != ctx->cfg.max_bounces) {  // the pending AOVs are those of the old setting (max_bounces < 1: no hit at all)
        int rc = ensureAovs(ctx);
        if (rc) return rc;
    }
    if (ctx->grouped() && cfg->accum_mode != ctx->cfg.accum_mode) ctx->combinedValid = false;
    ctx->cfg = *cfg;
    return SAIL_OK;
}

int sail_upload_scene(sail_ctx* ctx, const float* objects, int32_t n, const float* texparams, int32_t tn,
                      const float* lights, int32_t ln) {
    if (!ctx) return SAIL_ERR_INVALID;
    if (!ctx->mem[0].dScene) return setErr(ctx, SAIL_ERR_STATE, "context was not created successfully");
    char msg[256];
    msg[0] = 0;
    // packed into a scratch block: a rejected table must not disturb the scene the context is rendering
    SailSceneBlock* blk = new (std::nothrow) SailSceneBlock;
    if (!blk) return setErr(ctx, SAIL_ERR_NOMEM, "out of memory");
    int rc = sail_pack_scene(objects, n, texparams, tn, lights, ln, blk, msg, sizeof(msg));
    if (rc) { delete blk; return setErr(ctx, rc, "sail_upload_scene: %s", msg); }
    ctx->texparams.assign(texparams, texparams + (size_t)tn * SAIL_TEXPARAMS_LENGTH);
    ctx->lights.assign(lights, lights + (size_t)ln * SAIL_LIGHTS_LENGTH);
    ctx->tn = tn;
    ctx->ln = ln;
    rc = uploadBlock(ctx, *blk);
    delete blk;
    return rc;
}

int sail_upload_objects(sail_ctx* ctx, const float* objects, int32_t n) {
    if (!ctx) return SAIL_ERR_INVALID;
    if (!ctx->sceneLoaded) return setErr(ctx, SAIL_ERR_STATE, "sail_upload_objects before sail_upload_scene");
    char msg[256];
    msg[0] = 0;
    SailSceneBlock* blk = new (std::nothrow) SailSceneBlock;
    if (!blk) return setErr(ctx, SAIL_ERR_NOMEM, "out of memory");
    int rc = sail_pack_scene(objects, n, ctx->texparams.data(), ctx->tn, ctx->lights.data(), ctx->ln, blk, msg, sizeof(msg));
    if (rc) { delete blk; return setErr(ctx, rc, "sail_upload_objects: %s", msg); }
    rc = uploadBlock(ctx, *blk);
    delete blk;
    return rc;
}

int sail_scene_features(const float* objects, int32_t n, const float* texparams, int32_t tn, const float* lights,
                        int32_t ln, char* err, int32_t errlen) {
    SailSceneBlock* blk = new (std::nothrow) SailSceneBlock;
    if (!blk) return SAIL_ERR_NOMEM;
    int rc = sail_pack_scene(objects, n, texparams, tn, lights, ln, blk, err, errlen);
    int feat = blk->features;
    delete blk;
    return rc ? rc : feat;
}

int sail_jit_compile_check(const float* objects, int32_t n, const float* texparams, int32_t tn, const float* lights, int32_t ln,
                           int32_t kernel, char* info, int32_t infolen) {
    if (kernel != SAIL_KERNEL_MEGA && kernel != SAIL_KERNEL_WAVEFRONT && kernel != SAIL_KERNEL_QUEUE) return SAIL_ERR_INVALID;
    SailSceneBlock* blk = new (std::nothrow) SailSceneBlock;
    if (!blk) return SAIL_ERR_NOMEM;
    int rc = sail_pack_scene(objects, n, texparams, tn, lights, ln, blk, info, infolen);
    if (rc == 0) {
        SailJitKernel k;
        char why[400];
        why[0] = 0;
        const int F = blk->features & SAIL_F_ALL;
        SailJitJob* job = sail_jit_request(kernel == SAIL_KERNEL_QUEUE ? 2 : (kernel == SAIL_KERNEL_WAVEFRONT ? 1 : 0), F, SAIL_MIN_CTAS(F), 1, blk, why, sizeof(why));
        // SAIL_B200_JIT_CHECK_NOWAIT=1 (tests): only start the compilation, to exercise a host that exits while NVRTC is working
        const char* nowait = getenv("SAIL_B200_JIT_CHECK_NOWAIT");
        if (job && nowait && nowait[0] == '1') {
            if (info && infolen > 0) snprintf(info, (size_t)infolen, "compiling");
        } else if (job && sail_jit_poll(job, -1, 1, &k, why, sizeof(why)) == 0) {
            if (info && infolen > 0) snprintf(info, (size_t)infolen, "%s", k.info);
        } else {
            if (info && infolen > 0) snprintf(info, (size_t)infolen, "%s", why);
            rc = SAIL_ERR_STATE;
        }
    }
    delete blk;
    return rc;
}

int sail_set_row_bands(sail_ctx* ctx, int32_t band_rows, int32_t band_stride, int32_t band_first) {
    if (!ctx) return SAIL_ERR_INVALID;
    if (ctx->grouped()) return setErr(ctx, SAIL_ERR_STATE, "a sharded context assigns its row bands itself");
    if (band_stride < 1 || band_first < 0 || band_first >= band_stride || band_rows < SAIL_TILE_H || band_rows % SAIL_TILE_H)
        return setErr(ctx, SAIL_ERR_INVALID, "bad row bands: rows %d (multiple of %d), stride %d, first %d", band_rows, SAIL_TILE_H, band_stride, band_first);
    ctx->bandRows = band_rows;
    ctx->bandStride = band_stride;
    ctx->bandFirst = band_first;
    return SAIL_OK;
}

int sail_reset(sail_ctx* ctx) {
    if (!ctx || !ctx->mem[0].accum) return SAIL_ERR_INVALID;
    return resetAll(ctx);
}

int sail_render_n(sail_ctx* ctx, const sail_frame* frames, int32_t count) {
    if (!ctx || !frames) return SAIL_ERR_INVALID;
    if (!ctx->sceneLoaded) return setErr(ctx, SAIL_ERR_STATE, "sail_render before sail_upload_scene");
    // a sample-sharded context deals the frames out: every member still gets at most SAIL_MAX_FRAMES_PER_CALL of them
    const int maxCount = SAIL_MAX_FRAMES_PER_CALL * ((ctx->grouped() && ctx->shardMode == SAIL_SHARD_SAMPLE) ? ctx->groupSize : 1);
    if (count <= 0 || count > maxCount) return setErr(ctx, SAIL_ERR_INVALID, "frame count %d not in 1..%d", count, maxCount);
    // timeSinceStart is seconds since the renderer started (tracer.js:98): never negative.  The cosine sampler relies on
    // seed = timeSinceStart + bounce >= 1 (sail_shade.cuh cosineSampleHemisphere).
    for (int k = 0; k < count; k++)
        if (!(frames[k].timeSinceStart >= 0.0f)) return setErr(ctx, SAIL_ERR_INVALID, "frame %d: timeSinceStart must be >= 0", k);
    const bool sample = ctx->grouped() && ctx->shardMode == SAIL_SHARD_SAMPLE;
    const bool tile = ctx->grouped() && ctx->shardMode == SAIL_SHARD_TILE;
    int first = 0;
    if (sample && ctx->cfg.accum_mode == SAIL_ACCUM_MEAN) {
        // textureWeight = 0 restarts the running mean (mix(e, cache, 0) = e; renderer.js:57-60 resets sampleCount): the sums restart
        for (int k = count - 1; k >= 0; k--)
            if (frames[k].textureWeight == 0.0f) { first = k; break; }
        if (frames[first].textureWeight == 0.0f && ctx->sampleIndex + (uint64_t)first > 0) {
            const bool keepLast = ctx->haveLastFrame;
            int rc = resetAll(ctx);
            if (rc) return rc;
            ctx->haveLastFrame = keepLast;
        }
    }
    // the variant follows the scene's content, never the caller's mask hints
    const int F = ctx->hostScene.features & SAIL_F_ALL;
    const bool tryJit = ctx->cfg.specialise != SAIL_SPECIALISE_OFF;
    int kind = kernelKind(ctx, F, tryJit);
    const int small = (rngArgsSmall(ctx, frames, count) ? 1 : 0) | (ctx->cfg.rr_start_bounce > 0 ? 2 : 0);
    bool jit = false;
    if (tryJit) {
        jit = specialisedReady(ctx, kind, small, F, ctx->cfg.specialise == SAIL_SPECIALISE_REQUIRE);
        if (!jit && ctx->cfg.specialise == SAIL_SPECIALISE_REQUIRE)
            return setErr(ctx, SAIL_ERR_STATE, "scene-specialised kernel unavailable: %s", ctx->jitNote[kind][small]);
        if (jit) snprintf(ctx->kernelInfo, sizeof(ctx->kernelInfo), "%s", ctx->jitNote[kind][small]);
        else {
            const char* note = ctx->jitNote[kind][small];
            kind = kernelKind(ctx, F, false);  // the ahead-of-time variants have their own measured policy (no queue build among them)
            snprintf(ctx->kernelInfo, sizeof(ctx->kernelInfo), "aot %s<F=%d> (%s)", kind ? "wavefront" : "mega", F, note);
        }
    } else {
        snprintf(ctx->kernelInfo, sizeof(ctx->kernelInfo), "aot %s<F=%d> (specialise = off)", kind ? "wavefront" : "mega", F);
    }
    for (int i = 0; i < ctx->nlocal; i++) {
        SailMember& m = ctx->mem[i];
        SailLaunch L;
        L.W = ctx->W; L.H = ctx->H; L.y0 = ctx->y0; L.y1 = ctx->y1;
        L.bandRows = ctx->bandRows; L.bandStride = ctx->bandStride; L.bandFirst = ctx->bandFirst;
        if (tile) { L.bandRows = ctx->groupBandRows; L.bandStride = ctx->groupSize; L.bandFirst = m.index; }
        int tileRows = (ctx->y1 - ctx->y0 + SAIL_TILE_H - 1) / SAIL_TILE_H;
        if (L.bandStride > 1) {  // interleaved bands over the whole image
            int nbands = (ctx->H + L.bandRows - 1) / L.bandRows;
            int mine = (nbands - L.bandFirst + L.bandStride - 1) / L.bandStride;
            if (mine < 0) mine = 0;
            tileRows = mine * (L.bandRows / SAIL_TILE_H);
            L.y0 = 0; L.y1 = ctx->H;
        }
        L.maxBounces = ctx->cfg.max_bounces;
        L.accumMode = sample ? SAIL_ACCUM_SUM : ctx->cfg.accum_mode;
        L.rrStart = ctx->cfg.rr_start_bounce;
        L.rcpW = dm_rcp((float)ctx->W);
        L.rcpH = dm_rcp((float)ctx->H);
        L.accum = m.accum;
        L.scene = m.dScene;
        L.sceneSmem = (unsigned)sailSceneSmemBytes(ctx->hostScene.n, ctx->hostScene.ln);
        L.nframes = 0;
        for (int k = first; k < count; k++) {
            // sample sharding: frame number s (since the restart) belongs to member s mod N
            if (sample && (int)((ctx->sampleIndex + (uint64_t)(k - first)) % (uint64_t)ctx->groupSize) != m.index) continue;
            lowerFrame(frames[k], L.frames[L.nframes++]);
        }
        CU(cudaSetDevice(m.device));
        CU(cudaEventRecord(m.evStart, m.stream));
        if (L.nframes > 0 && tileRows > 0) {
            dim3 grid((ctx->W + SAIL_TILE_W - 1) / SAIL_TILE_W, tileRows);
            if (jit && kind == 2) {
                int qrc = launchQueue(ctx, m, small, L.frames, L.nframes, L.accumMode);
                if (qrc) return qrc;
            } else if (jit) {
                char why[200];
                unsigned gx = grid.x;
                if (kind == 1) gx = ((unsigned)ctx->W + m.jitBlock[1][small] / 8u - 1u) / (m.jitBlock[1][small] / 8u);  // tile = (threads / 8) x 8
                if (sail_jit_launch(m.jitFn[kind][small], &L, gx, grid.y, m.jitBlock[kind][small], L.sceneSmem, m.stream, why, sizeof(why)) != 0)
                    return setErr(ctx, SAIL_ERR_CUDA, "%s", why);
            } else if (kind == 1) CU(launchWave(F, L, grid, m.stream));
            else CU(launchMega(F, L, grid, m.stream));
            if (!(jit && kind == 2)) ctx->launches++;
        }
        CU(cudaEventRecord(m.evStop, m.stream));
        m.timed = true;
    }
    if (sample) ctx->sampleIndex += (uint64_t)(count - first);
    ctx->lastFrame = frames[count - 1];
    ctx->haveLastFrame = true;
    ctx->combinedValid = false;
    ctx->aovValid = false;
    return SAIL_OK;
}

int sail_render(sail_ctx* ctx, const float matrix_colmajor[16], const float eye[3], float textureWeight, float timeSinceStart) {
    if (!ctx || !matrix_colmajor || !eye) return SAIL_ERR_INVALID;
    sail_frame f;
    memcpy(f.matrix, matrix_colmajor, sizeof(f.matrix));
    memcpy(f.eye, eye, sizeof(f.eye));
    f.textureWeight = textureWeight;
    f.timeSinceStart = timeSinceStart;
    return sail_render_n(ctx, &f, 1);
}

int sail_specialise_wait(sail_ctx* ctx) {
    if (!ctx) return SAIL_ERR_INVALID;
    if (!ctx->sceneLoaded) return setErr(ctx, SAIL_ERR_STATE, "sail_specialise_wait before sail_upload_scene");
    if (ctx->cfg.specialise == SAIL_SPECIALISE_OFF) return 1;
    const int F = ctx->hostScene.features & SAIL_F_ALL;
    const int kind = kernelKind(ctx, F, true);
    // the variant the last frames asked for (a fresh context: the one a session starts with)
    const int small = (ctx->haveLastFrame ? (rngArgsSmall(ctx, &ctx->lastFrame, 1) ? 1 : 0) : 1) | (ctx->cfg.rr_start_bounce > 0 ? 2 : 0);
    return specialisedReady(ctx, kind, small, F, true) ? 0 : 1;
}

int sail_synchronize(sail_ctx* ctx) {
    if (!ctx || !ctx->mem[0].stream) return SAIL_ERR_INVALID;
    for (int i = 0; i < ctx->nlocal; i++) {
        CU(cudaSetDevice(ctx->mem[i].device));
        CU(cudaStreamSynchronize(ctx->mem[i].stream));
    }
    return SAIL_OK;
}

static void* rawBuffer(sail_ctx* ctx, int32_t which) {
    SailMember& r = ctx->root();
    switch (which) {
        case SAIL_AOV_COLOR: return r.accum;
        case SAIL_AOV_NORMAL: return r.normal;
        case SAIL_AOV_POSITION: return r.position;
        case SAIL_AOV_HIT_ID: return r.hitId;
        case SAIL_AOV_DEPTH: return r.depth;
    }
    return nullptr;
}
void* sail_device_ptr(sail_ctx* ctx, int32_t which) {
    if (!ctx) return nullptr;
    if (which != SAIL_AOV_COLOR && ensureAovs(ctx) != SAIL_OK) return nullptr;  // queued on the context's stream
    return rawBuffer(ctx, which);
}
void* sail_stream(sail_ctx* ctx) { return ctx ? (void*)ctx->root().stream : nullptr; }

int sail_read_aov(sail_ctx* ctx, int32_t which, void* out, size_t out_bytes) {
    if (!ctx) return SAIL_ERR_INVALID;
    const void* src = rawBuffer(ctx, which);
    if (!src) return setErr(ctx, SAIL_ERR_INVALID, "unknown AOV %d", which);
    size_t bytes = npix(ctx) * ((which == SAIL_AOV_HIT_ID || which == SAIL_AOV_DEPTH) ? 4 : 12);
    if (which == SAIL_AOV_COLOR) {
        const float* img = nullptr;
        int rc = ensureCombined(ctx, &img);  // collective on a joined context
        if (rc) return rc;
        src = img;
    } else {
        int rc = ensureAovs(ctx);
        if (rc) return rc;
    }
    if (!ctx->isRoot()) return SAIL_OK;  // images arrive on the root
    if (!out) return SAIL_ERR_INVALID;
    if (out_bytes != bytes) return setErr(ctx, SAIL_ERR_INVALID, "AOV %d needs %zu bytes, caller gave %zu", which, bytes, out_bytes);
    SailMember& r = ctx->root();
    CU(cudaSetDevice(r.device));
    CU(cudaMemcpyAsync(out, src, bytes, cudaMemcpyDeviceToHost, r.stream));
    if (which == SAIL_AOV_COLOR) { int rc = markConsumed(ctx); if (rc) return rc; }
    CU(cudaStreamSynchronize(r.stream));
    return SAIL_OK;
}
int sail_read_accum(sail_ctx* ctx, float* rgb, size_t out_bytes) { return sail_read_aov(ctx, SAIL_AOV_COLOR, rgb, out_bytes); }

int sail_write_accum(sail_ctx* ctx, const float* rgb, size_t in_bytes) {
    if (!ctx || !rgb) return SAIL_ERR_INVALID;
    size_t bytes = npix(ctx) * 12;
    if (in_bytes != bytes) return setErr(ctx, SAIL_ERR_INVALID, "accumulation buffer is %zu bytes, caller gave %zu", bytes, in_bytes);
    if (ctx->grouped()) {
        // the image becomes the root's share, as one sample; the other members start from zero
        const sail_frame keep = ctx->lastFrame;
        const bool had = ctx->haveLastFrame;
        int rc = resetAll(ctx);
        if (rc) return rc;
        ctx->lastFrame = keep;
        ctx->haveLastFrame = had;
        ctx->sampleIndex = 1;
        if (!ctx->isRoot()) return SAIL_OK;
    }
    SailMember& r = ctx->root();
    CU(cudaSetDevice(r.device));
    CU(cudaMemcpyAsync(r.accum, rgb, bytes, cudaMemcpyHostToDevice, r.stream));
    CU(cudaStreamSynchronize(r.stream));
    return SAIL_OK;
}

int sail_scale_accum(sail_ctx* ctx, float s) {
    if (!ctx || !ctx->mem[0].accum) return SAIL_ERR_INVALID;
    size_t n = npix(ctx) * 3;
    for (int i = 0; i < ctx->nlocal; i++) {
        SailMember& m = ctx->mem[i];
        CU(cudaSetDevice(m.device));
        sail_scale_kernel<<<(unsigned)((n + 255) / 256), 256, 0, m.stream>>>(m.accum, n, s);
        ctx->launches++;
        CU(cudaGetLastError());
    }
    ctx->combinedValid = false;
    return SAIL_OK;
}

int sail_present_begin(sail_ctx* ctx, uint8_t* rgba8) {
    if (!ctx || !ctx->mem[0].accum) return SAIL_ERR_INVALID;
    if (ctx->isRoot() && !rgba8) return SAIL_ERR_INVALID;
    if (ctx->copyInFlight) return setErr(ctx, SAIL_ERR_STATE, "sail_present_begin: the previous image has not been collected with sail_present_end");
    const float* img = nullptr;
    int rc = ensureCombined(ctx, &img);
    if (rc) return rc;
    if (!ctx->isRoot()) { ctx->copyInFlight = true; return SAIL_OK; }
    if (filterNeedsAovs(ctx->cfg.filter_id) && (rc = ensureAovs(ctx)) != 0) return rc;
    SailMember& r = ctx->root();
    CU(cudaSetDevice(r.device));
    size_t n = npix(ctx);
    if (!ctx->presentU8) CU(cudaMalloc(&ctx->presentU8, n * 4));
    PresentParams P;
    fillPresentParams(ctx, P, img);
    P.out_u8 = ctx->presentU8;
    // the filter overwrites the device image: it must not overtake the previous copy (long finished in practice)
    if (ctx->copiedOnce) CU(cudaStreamWaitEvent(r.stream, ctx->evCopied, 0));
    if (sailPresentWantsPadded(P.filter_id)) {
        if (!ctx->presentPad) CU(cudaMalloc(&ctx->presentPad, n * sizeof(float4)));
        ctx->launches++;  // sail_pad_rgb_kernel
    }
    if (P.filter_id == 9) {  // wavelet: the position map as well
        if (!ctx->presentPad2) CU(cudaMalloc(&ctx->presentPad2, n * sizeof(float4)));
        ctx->launches++;
    }
    CU(sailLaunchPresent(P, ctx->presentPad, ctx->presentPad2, r.stream));
    ctx->launches++;
    CU(cudaGetLastError());
    CU(cudaEventRecord(ctx->evPresented, r.stream));
    if ((rc = markConsumed(ctx)) != 0) return rc;
    CU(cudaStreamWaitEvent(ctx->copyStream, ctx->evPresented, 0));
    CU(cudaMemcpyAsync(rgba8, ctx->presentU8, n * 4, cudaMemcpyDeviceToHost, ctx->copyStream));
    CU(cudaEventRecord(ctx->evCopied, ctx->copyStream));
    ctx->copyInFlight = true;
    ctx->copiedOnce = true;
    return SAIL_OK;
}

int sail_present_end(sail_ctx* ctx) {
    if (!ctx) return SAIL_ERR_INVALID;
    if (!ctx->copyInFlight) return setErr(ctx, SAIL_ERR_STATE, "sail_present_end without sail_present_begin");
    ctx->copyInFlight = false;
    if (!ctx->isRoot()) return SAIL_OK;
    CU(cudaSetDevice(ctx->root().device));
    CU(cudaEventSynchronize(ctx->evCopied));
    return SAIL_OK;
}

int sail_present(sail_ctx* ctx, float* rgba_f32, uint8_t* rgba8) {
    if (!ctx || !ctx->mem[0].accum) return SAIL_ERR_INVALID;
    if (rgba8 && ctx->copyInFlight) return setErr(ctx, SAIL_ERR_STATE, "sail_present while a sail_present_begin image is in flight");
    const float* img = nullptr;
    int rc = ensureCombined(ctx, &img);
    if (rc) return rc;
    if (!ctx->isRoot()) return SAIL_OK;
    if (filterNeedsAovs(ctx->cfg.filter_id) && (rc = ensureAovs(ctx)) != 0) return rc;
    SailMember& r = ctx->root();
    CU(cudaSetDevice(r.device));
    size_t n = npix(ctx);
    if (rgba_f32 && !ctx->presentF32) CU(cudaMalloc(&ctx->presentF32, n * 4 * sizeof(float)));
    if (rgba8 && !ctx->presentU8) CU(cudaMalloc(&ctx->presentU8, n * 4));
    PresentParams P;
    fillPresentParams(ctx, P, img);
    P.out_f32 = rgba_f32 ? ctx->presentF32 : nullptr;
    P.out_u8 = rgba8 ? ctx->presentU8 : nullptr;
    if (sailPresentWantsPadded(P.filter_id)) {
        if (!ctx->presentPad) CU(cudaMalloc(&ctx->presentPad, n * sizeof(float4)));
        ctx->launches++;  // sail_pad_rgb_kernel
    }
    if (P.filter_id == 9) {  // wavelet: the position map as well
        if (!ctx->presentPad2) CU(cudaMalloc(&ctx->presentPad2, n * sizeof(float4)));
        ctx->launches++;
    }
    CU(sailLaunchPresent(P, ctx->presentPad, ctx->presentPad2, r.stream));
    ctx->launches++;
    CU(cudaGetLastError());
    if ((rc = markConsumed(ctx)) != 0) return rc;
    if (rgba_f32) CU(cudaMemcpyAsync(rgba_f32, ctx->presentF32, n * 4 * sizeof(float), cudaMemcpyDeviceToHost, r.stream));
    if (rgba8) CU(cudaMemcpyAsync(rgba8, ctx->presentU8, n * 4, cudaMemcpyDeviceToHost, r.stream));
    CU(cudaStreamSynchronize(r.stream));
    return SAIL_OK;
}

int sail_intersect_rays(sail_ctx* ctx, const float* rays, int32_t count, int32_t* hit_id, float* t, float* normal, int32_t* mat_category) {
    if (!ctx || !rays || count < 0) return SAIL_ERR_INVALID;
    if (!ctx->sceneLoaded) return setErr(ctx, SAIL_ERR_STATE, "sail_intersect_rays before sail_upload_scene");
    if (count == 0) return SAIL_OK;
    SailMember& m = ctx->root();
    CU(cudaSetDevice(m.device));
    float *dR = nullptr, *dT = nullptr, *dN = nullptr;
    int *dI = nullptr, *dM = nullptr;
    size_t c = (size_t)count;
    CU(cudaMalloc(&dR, c * 6 * sizeof(float)));
    CU(cudaMalloc(&dT, c * sizeof(float)));
    CU(cudaMalloc(&dN, c * 3 * sizeof(float)));
    CU(cudaMalloc(&dI, c * sizeof(int)));
    CU(cudaMalloc(&dM, c * sizeof(int)));
    CU(cudaMemcpyAsync(dR, rays, c * 6 * sizeof(float), cudaMemcpyHostToDevice, m.stream));
    sail_intersect_kernel<<<(unsigned)((c + SAIL_CTA_THREADS - 1) / SAIL_CTA_THREADS), SAIL_CTA_THREADS, sailSceneSmemBytes(ctx->hostScene.n, ctx->hostScene.ln), m.stream>>>(m.dScene, dR, count, dI, dT, dN, dM);
    ctx->launches++;
    CU(cudaGetLastError());
    if (hit_id) CU(cudaMemcpyAsync(hit_id, dI, c * sizeof(int), cudaMemcpyDeviceToHost, m.stream));
    if (t) CU(cudaMemcpyAsync(t, dT, c * sizeof(float), cudaMemcpyDeviceToHost, m.stream));
    if (normal) CU(cudaMemcpyAsync(normal, dN, c * 3 * sizeof(float), cudaMemcpyDeviceToHost, m.stream));
    if (mat_category) CU(cudaMemcpyAsync(mat_category, dM, c * sizeof(int), cudaMemcpyDeviceToHost, m.stream));
    CU(cudaStreamSynchronize(m.stream));
    cudaFree(dR); cudaFree(dT); cudaFree(dN); cudaFree(dI); cudaFree(dM);
    return SAIL_OK;
}

// ---- sail_intersect_rays_noalloc -------------------------------------------------------------------------------------
// The scratch of sail_intersect_rays() (device copies of the rays and of the four outputs, cudaMalloc'ed and freed per call)
// as ONE file-scope device array, for callers that must not allocate: 12 words per ray.  Every GPU that loads the library
// holds these SAIL_NOALLOC_MAX_RAYS * 48 bytes (3 MiB) until the process exits.  All calls on a device share the array and it
// keeps the last call's contents; nothing in it has to start at zero (the kernel writes every output element below `count`
// before the copies read it), but two calls that overlap on the GPU (different contexts / streams) race on it.
static __device__ __align__(256) float sail_noalloc_scratch[(size_t)SAIL_NOALLOC_MAX_RAYS * 12];

namespace {
// In host code `sail_noalloc_scratch` is only a handle for the runtime, not a device address: the address is asked for
// once per device, with that device current.
float* noallocScratch(int device, cudaError_t* e) {
    static float* ptr[64] = {nullptr};
    static std::mutex m;
    std::lock_guard<std::mutex> lock(m);
    *e = cudaSuccess;
    if (device < 0 || device >= 64) { *e = cudaErrorInvalidDevice; return nullptr; }
    if (!ptr[device]) {
        void* p = nullptr;
        *e = cudaGetSymbolAddress(&p, sail_noalloc_scratch);  // the caller has made `device` current
        if (*e != cudaSuccess) return nullptr;
        ptr[device] = static_cast<float*>(p);
    }
    return ptr[device];
}
}  // namespace

int sail_intersect_rays_noalloc(sail_ctx* ctx, const float* rays, int32_t count, int32_t* hit_id, float* t, float* normal, int32_t* mat_category) {
    if (!ctx || !rays || count < 0) return SAIL_ERR_INVALID;
    if (!ctx->sceneLoaded) return setErr(ctx, SAIL_ERR_STATE, "sail_intersect_rays_noalloc before sail_upload_scene");
    if (count > SAIL_NOALLOC_MAX_RAYS)
        return setErr(ctx, SAIL_ERR_INVALID, "sail_intersect_rays_noalloc: %d rays exceed the fixed scratch of SAIL_NOALLOC_MAX_RAYS = %d; split the batch or use sail_intersect_rays",
                      count, SAIL_NOALLOC_MAX_RAYS);
    if (count == 0) return SAIL_OK;
    SailMember& m = ctx->root();
    CU(cudaSetDevice(m.device));
    cudaError_t se;
    float* base = noallocScratch(m.device, &se);
    if (!base) return setErr(ctx, SAIL_ERR_CUDA, "cudaGetSymbolAddress(sail_noalloc_scratch) failed: %s", cudaGetErrorString(se));
    const size_t c = (size_t)count, cap = (size_t)SAIL_NOALLOC_MAX_RAYS;
    float* dR = base;                                         // cap x 6
    float* dT = base + cap * 6;                               // cap
    float* dN = base + cap * 7;                               // cap x 3
    int* dI = reinterpret_cast<int*>(base + cap * 10);        // cap
    int* dM = reinterpret_cast<int*>(base + cap * 11);        // cap
    CU(cudaMemcpyAsync(dR, rays, c * 6 * sizeof(float), cudaMemcpyHostToDevice, m.stream));
    sail_intersect_kernel<<<(unsigned)((c + SAIL_CTA_THREADS - 1) / SAIL_CTA_THREADS), SAIL_CTA_THREADS, sailSceneSmemBytes(ctx->hostScene.n, ctx->hostScene.ln), m.stream>>>(m.dScene, dR, count, dI, dT, dN, dM);
    ctx->launches++;
    CU(cudaGetLastError());
    if (hit_id) CU(cudaMemcpyAsync(hit_id, dI, c * sizeof(int), cudaMemcpyDeviceToHost, m.stream));
    if (t) CU(cudaMemcpyAsync(t, dT, c * sizeof(float), cudaMemcpyDeviceToHost, m.stream));
    if (normal) CU(cudaMemcpyAsync(normal, dN, c * 3 * sizeof(float), cudaMemcpyDeviceToHost, m.stream));
    if (mat_category) CU(cudaMemcpyAsync(mat_category, dM, c * sizeof(int), cudaMemcpyDeviceToHost, m.stream));
    CU(cudaStreamSynchronize(m.stream));
    return SAIL_OK;
}

int sail_trace_dump(sail_ctx* ctx, const sail_frame* frame, int32_t* hit_id, float* t, float* rays, float* radiance) {
    if (!ctx || !frame) return SAIL_ERR_INVALID;
    if (!ctx->sceneLoaded) return setErr(ctx, SAIL_ERR_STATE, "sail_trace_dump before sail_upload_scene");
    SailMember& m = ctx->root();
    CU(cudaSetDevice(m.device));
    size_t n = npix(ctx), B = (size_t)ctx->cfg.max_bounces;
    if (B == 0) return setErr(ctx, SAIL_ERR_INVALID, "max_bounces is 0");
    SailDumpLaunch L;
    L.W = ctx->W; L.H = ctx->H; L.maxBounces = (int)B; L.rrStart = ctx->cfg.rr_start_bounce;
    L.rcpW = dm_rcp((float)ctx->W); L.rcpH = dm_rcp((float)ctx->H);
    L.scene = m.dScene;
    L.hitId = nullptr; L.t = nullptr; L.rays = nullptr; L.radiance = nullptr;
    if (!(frame->timeSinceStart >= 0.0f)) return setErr(ctx, SAIL_ERR_INVALID, "timeSinceStart must be >= 0");
    lowerFrame(*frame, L.frame);
    if (hit_id) CU(cudaMalloc(&L.hitId, n * B * sizeof(int)));
    if (t) CU(cudaMalloc(&L.t, n * B * sizeof(float)));
    if (rays) CU(cudaMalloc(&L.rays, n * B * 6 * sizeof(float)));
    if (radiance) CU(cudaMalloc(&L.radiance, n * 3 * sizeof(float)));
    dim3 grid((ctx->W + SAIL_TILE_W - 1) / SAIL_TILE_W, (ctx->H + SAIL_TILE_H - 1) / SAIL_TILE_H);
    sail_trace_dump_kernel<<<grid, SAIL_CTA_THREADS, sailSceneSmemBytes(ctx->hostScene.n, ctx->hostScene.ln), m.stream>>>(L);
    ctx->launches++;
    CU(cudaGetLastError());
    if (hit_id) CU(cudaMemcpyAsync(hit_id, L.hitId, n * B * sizeof(int), cudaMemcpyDeviceToHost, m.stream));
    if (t) CU(cudaMemcpyAsync(t, L.t, n * B * sizeof(float), cudaMemcpyDeviceToHost, m.stream));
    if (rays) CU(cudaMemcpyAsync(rays, L.rays, n * B * 6 * sizeof(float), cudaMemcpyDeviceToHost, m.stream));
    if (radiance) CU(cudaMemcpyAsync(radiance, L.radiance, n * 3 * sizeof(float), cudaMemcpyDeviceToHost, m.stream));
    CU(cudaStreamSynchronize(m.stream));
    cudaFree(L.hitId); cudaFree(L.t); cudaFree(L.rays); cudaFree(L.radiance);
    return SAIL_OK;
}

int sail_eval_math(sail_ctx* ctx, int32_t which, const float* a, const float* b, float* out, int32_t count) {
    if (!ctx || !a || !out || count < 0 || which < 0 || which > 12) return SAIL_ERR_INVALID;
    if (count == 0) return SAIL_OK;
    SailMember& m = ctx->root();
    CU(cudaSetDevice(m.device));
    float *dA = nullptr, *dB = nullptr, *dO = nullptr;
    size_t bytes = (size_t)count * sizeof(float);
    CU(cudaMalloc(&dA, bytes));
    CU(cudaMalloc(&dO, bytes));
    CU(cudaMemcpyAsync(dA, a, bytes, cudaMemcpyHostToDevice, m.stream));
    if (b) {
        CU(cudaMalloc(&dB, bytes));
        CU(cudaMemcpyAsync(dB, b, bytes, cudaMemcpyHostToDevice, m.stream));
    }
    sail_eval_math_kernel<<<(count + 255) / 256, 256, 0, m.stream>>>(which, dA, dB, dO, count);
    ctx->launches++;
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(out, dO, bytes, cudaMemcpyDeviceToHost, m.stream));
    CU(cudaStreamSynchronize(m.stream));
    cudaFree(dA); cudaFree(dB); cudaFree(dO);
    return SAIL_OK;
}

int sail_selftest_math(sail_ctx* ctx, uint64_t* mismatches) {
    if (!ctx || !mismatches) return SAIL_ERR_INVALID;
    SailMember& m = ctx->root();
    CU(cudaSetDevice(m.device));
    unsigned long long* d = nullptr;
    CU(cudaMalloc(&d, sizeof(unsigned long long)));
    CU(cudaMemsetAsync(d, 0, sizeof(unsigned long long), m.stream));
    sail_selftest_rcp_kernel<<<(1u << 20) / 256, 256, 0, m.stream>>>(d);
    sail_selftest_unit_kernel<<<(1u << 19) / 256, 256, 0, m.stream>>>(d);
    ctx->launches++;
    CU(cudaGetLastError());
    unsigned long long h = 0;
    CU(cudaMemcpyAsync(&h, d, sizeof(h), cudaMemcpyDeviceToHost, m.stream));
    CU(cudaStreamSynchronize(m.stream));
    cudaFree(d);
    *mismatches = h;
    return SAIL_OK;
}

uint64_t sail_launch_count(const sail_ctx* ctx) { return ctx ? ctx->launches : 0; }

int sail_last_render_ms(sail_ctx* ctx, float* ms) {
    if (!ctx || !ms) return SAIL_ERR_INVALID;
    if (!ctx->mem[0].timed) return setErr(ctx, SAIL_ERR_STATE, "no render has been issued yet");
    float worst = 0.0f;
    for (int i = 0; i < ctx->nlocal; i++) {  // the slowest member
        SailMember& m = ctx->mem[i];
        float t = 0.0f;
        CU(cudaSetDevice(m.device));
        CU(cudaEventSynchronize(m.evStop));
        CU(cudaEventElapsedTime(&t, m.evStart, m.evStop));
        if (t > worst) worst = t;
    }
    *ms = worst;
    return SAIL_OK;
}

int sail_last_combine_ms(sail_ctx* ctx, float* ms) {
    if (!ctx || !ms) return SAIL_ERR_INVALID;
    if (!ctx->grouped() || !ctx->isRoot() || !ctx->combineTimed) return setErr(ctx, SAIL_ERR_STATE, "no combine has run on this context (sharded contexts, root only)");
    CU(cudaSetDevice(ctx->root().device));
    CU(cudaEventSynchronize(ctx->evCombStop));
    CU(cudaEventElapsedTime(ms, ctx->evCombStart, ctx->evCombStop));
    return SAIL_OK;
}

int sail_kernel_info(sail_ctx* ctx, char* buf, int32_t len) {
    if (!ctx || !buf || len < 1) return SAIL_ERR_INVALID;
    return snprintf(buf, (size_t)len, "%s", ctx->kernelInfo[0] ? ctx->kernelInfo : "no trace kernel launched yet");
}

void* sail_alloc_host(uint64_t bytes) {
    void* p = nullptr;
    if (bytes == 0 || cudaHostAlloc(&p, (size_t)bytes, cudaHostAllocPortable) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    return p;
}
void sail_free_host(void* p) {
    if (p) cudaFreeHost(p);
}

}  // extern "C"
