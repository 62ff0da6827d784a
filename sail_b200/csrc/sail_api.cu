// sail_api.cu -- libsail_b200.so: the C ABI of include/sail_b200.h over the sm_100a kernels.
//
// Stands in for the reference's render orchestration + GPU abstraction on the hot path:
// Renderer/Tracer (/root/reference/src/core/renderer.js:8-73, src/core/tracer.js:8-101) and ShaderProgram's
// framebuffer/texture/uniform plumbing (src/core/webgl.js:4-135).  Device-resident FP32 buffers replace
// ShaderProgram.frameCache[0..3]; a packed scene block replaces the three R32F textures; CUDA kernels
// replace the trace and render programs.  No CPU fallback exists in this file or anywhere in the library.
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <new>
#include <vector>

#include "../../include/sail_b200.h"
#include "sail_diag.cuh"
#include "sail_jit.h"
#include "sail_mega.cuh"
#include "sail_present.cuh"

struct sail_ctx {
    int device = 0;
    int W = 0, H = 0, y0 = 0, y1 = 0;
    int bandRows = 0, bandStride = 1, bandFirst = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t evStart = nullptr, evStop = nullptr;
    // double-buffered display (sail_present_begin / _end): copy stream, filter-done and copy-done events
    cudaStream_t copyStream = nullptr;
    cudaEvent_t evPresented = nullptr, evCopied = nullptr;
    bool copyInFlight = false, copiedOnce = false;
    bool timed = false;
    sail_config cfg;
    bool sceneLoaded = false;
    std::vector<float> texparams, lights;  // kept for sail_upload_objects (tracer.js:25-40 re-uses them)
    int tn = 0, ln = 0;
    SailSceneBlock hostScene;
    SailSceneBlock* dScene = nullptr;
    float *accum = nullptr, *normal = nullptr, *position = nullptr, *depth = nullptr;
    int* hitId = nullptr;
    float* presentF32 = nullptr;
    unsigned char* presentU8 = nullptr;
    uint64_t launches = 0;
    // scene-specialised kernels (sail_jit.cpp): resolved lazily per kernel kind, dropped when the scene changes
    void* jitFn[3] = {nullptr, nullptr, nullptr};  // kind 0 mega, 1 wavefront, 2 server
    bool jitTried[3] = {false, false, false};
    unsigned jitBlock[3] = {0, 0, 0};
    int jitPathWarps[3] = {0, 0, 0};
    char jitNote[3][320];
    char kernelInfo[400];
    char err[512];
    sail_ctx() { err[0] = 0; kernelInfo[0] = 0; jitNote[0][0] = jitNote[1][0] = jitNote[2][0] = 0; }
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

// A scene the server-warp kernel (sail_server.cuh) is meant for: cubes / Cornell boxes plus other shapes.
bool hasBoxesAndOthers(const SailSceneBlock& b) {
    bool box = false, other = false;
    for (int i = 0; i < b.n; i++) {
        const bool isBox = b.shape[i].category == SAIL_CUBE || b.shape[i].category == SAIL_CORNELLBOX;
        box |= isBox;
        other |= !isBox;
    }
    return box && other;
}
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

int uploadBlock(sail_ctx* ctx) {
    CU(cudaSetDevice(ctx->device));
    CU(cudaMemcpyAsync(ctx->dScene, &ctx->hostScene, sizeof(SailSceneBlock), cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));  // hostScene may be rewritten by the next upload
    ctx->sceneLoaded = true;
    ctx->jitTried[0] = ctx->jitTried[1] = ctx->jitTried[2] = false;  // the structure may have changed: look the specialised kernels up again
    ctx->jitFn[0] = ctx->jitFn[1] = ctx->jitFn[2] = nullptr;
    return SAIL_OK;
}

}  // namespace

extern "C" {

const char* sail_version(void) { return "sail-b200 0.1 (sm_100a)"; }
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
    if (info->device < 0 || info->device >= ndev) return setErr(ctx, SAIL_ERR_INVALID, "device %d out of range (%d devices)", info->device, ndev);
    ctx->device = info->device;
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, ctx->device));
    if (prop.major != 10)
        return setErr(ctx, SAIL_ERR_CUDA, "device %d is sm_%d%d; this library carries sm_100a code only", ctx->device, prop.major, prop.minor);
    CU(cudaSetDevice(ctx->device));
    CU(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    CU(cudaEventCreate(&ctx->evStart));
    CU(cudaEventCreate(&ctx->evStop));
    CU(cudaStreamCreateWithFlags(&ctx->copyStream, cudaStreamNonBlocking));
    CU(cudaEventCreateWithFlags(&ctx->evPresented, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&ctx->evCopied, cudaEventDisableTiming));
    size_t n = npix(ctx);
    CU(cudaMalloc(&ctx->dScene, sizeof(SailSceneBlock)));
    CU(cudaMalloc(&ctx->accum, n * 3 * sizeof(float)));
    CU(cudaMalloc(&ctx->normal, n * 3 * sizeof(float)));
    CU(cudaMalloc(&ctx->position, n * 3 * sizeof(float)));
    CU(cudaMalloc(&ctx->depth, n * sizeof(float)));
    CU(cudaMalloc(&ctx->hitId, n * sizeof(int)));
    sail_default_config(&ctx->cfg);
    return sail_reset(ctx);
}

void sail_destroy(sail_ctx* ctx) {
    if (!ctx) return;
    if (ctx->stream) {
        cudaSetDevice(ctx->device);
        cudaStreamSynchronize(ctx->stream);
        if (ctx->copyStream) cudaStreamSynchronize(ctx->copyStream);
    }
    cudaFree(ctx->dScene); cudaFree(ctx->accum); cudaFree(ctx->normal); cudaFree(ctx->position);
    cudaFree(ctx->depth); cudaFree(ctx->hitId); cudaFree(ctx->presentF32); cudaFree(ctx->presentU8);
    if (ctx->evStart) cudaEventDestroy(ctx->evStart);
    if (ctx->evStop) cudaEventDestroy(ctx->evStop);
    if (ctx->evPresented) cudaEventDestroy(ctx->evPresented);
    if (ctx->evCopied) cudaEventDestroy(ctx->evCopied);
    if (ctx->copyStream) cudaStreamDestroy(ctx->copyStream);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

const char* sail_last_error(const sail_ctx* ctx) { return ctx ? ctx->err : "null context"; }

int sail_configure(sail_ctx* ctx, const sail_config* cfg) {
    if (!ctx || !cfg) return SAIL_ERR_INVALID;
    if (cfg->trace_id != 0) return setErr(ctx, SAIL_ERR_INVALID, "unknown trace plugin id %d (only 0 = path)", cfg->trace_id);
    if (cfg->max_bounces < 0 || cfg->max_bounces > 1024) return setErr(ctx, SAIL_ERR_INVALID, "max_bounces %d out of range", cfg->max_bounces);
    if (cfg->filter_id < 0 || cfg->filter_id > SAIL_FILTER_TONEMAPPING) return setErr(ctx, SAIL_ERR_INVALID, "unknown filter id %d", cfg->filter_id);
    if (cfg->accum_mode != SAIL_ACCUM_MEAN && cfg->accum_mode != SAIL_ACCUM_SUM) return setErr(ctx, SAIL_ERR_INVALID, "unknown accum_mode %d", cfg->accum_mode);
    if (cfg->kernel < SAIL_KERNEL_MEGA || cfg->kernel > SAIL_KERNEL_SERVER) return setErr(ctx, SAIL_ERR_INVALID, "unknown kernel %d", cfg->kernel);
    if (cfg->specialise < SAIL_SPECIALISE_AUTO || cfg->specialise > SAIL_SPECIALISE_REQUIRE) return setErr(ctx, SAIL_ERR_INVALID, "unknown specialise mode %d", cfg->specialise);
    ctx->cfg = *cfg;
    return SAIL_OK;
}

int sail_upload_scene(sail_ctx* ctx, const float* objects, int32_t n, const float* texparams, int32_t tn,
                      const float* lights, int32_t ln) {
    if (!ctx) return SAIL_ERR_INVALID;
    if (!ctx->dScene) return setErr(ctx, SAIL_ERR_STATE, "context was not created successfully");
    char msg[256];
    msg[0] = 0;
    int rc = sail_pack_scene(objects, n, texparams, tn, lights, ln, &ctx->hostScene, msg, sizeof(msg));
    if (rc) return setErr(ctx, rc, "sail_upload_scene: %s", msg);
    ctx->texparams.assign(texparams, texparams + (size_t)tn * SAIL_TEXPARAMS_LENGTH);
    ctx->lights.assign(lights, lights + (size_t)ln * SAIL_LIGHTS_LENGTH);
    ctx->tn = tn;
    ctx->ln = ln;
    return uploadBlock(ctx);
}

int sail_upload_objects(sail_ctx* ctx, const float* objects, int32_t n) {
    if (!ctx) return SAIL_ERR_INVALID;
    if (!ctx->sceneLoaded) return setErr(ctx, SAIL_ERR_STATE, "sail_upload_objects before sail_upload_scene");
    char msg[256];
    msg[0] = 0;
    int rc = sail_pack_scene(objects, n, ctx->texparams.data(), ctx->tn, ctx->lights.data(), ctx->ln, &ctx->hostScene, msg, sizeof(msg));
    if (rc) return setErr(ctx, rc, "sail_upload_objects: %s", msg);
    return uploadBlock(ctx);
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
    if (kernel != SAIL_KERNEL_MEGA && kernel != SAIL_KERNEL_WAVEFRONT && kernel != SAIL_KERNEL_SERVER) return SAIL_ERR_INVALID;
    SailSceneBlock* blk = new (std::nothrow) SailSceneBlock;
    if (!blk) return SAIL_ERR_NOMEM;
    int rc = sail_pack_scene(objects, n, texparams, tn, lights, ln, blk, info, infolen);
    if (rc == 0) {
        SailJitKernel k;
        char why[400];
        why[0] = 0;
        const int F = blk->features & SAIL_F_ALL;
        if (sail_jit_get(-1, kernel == SAIL_KERNEL_SERVER ? 2 : (kernel == SAIL_KERNEL_WAVEFRONT ? 1 : 0), F, SAIL_MIN_CTAS(F), blk, &k, why, sizeof(why)) == 0) {
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
    if (band_stride < 1 || band_first < 0 || band_first >= band_stride || band_rows < SAIL_TILE_H || band_rows % SAIL_TILE_H)
        return setErr(ctx, SAIL_ERR_INVALID, "bad row bands: rows %d (multiple of %d), stride %d, first %d", band_rows, SAIL_TILE_H, band_stride, band_first);
    ctx->bandRows = band_rows;
    ctx->bandStride = band_stride;
    ctx->bandFirst = band_first;
    return SAIL_OK;
}

int sail_reset(sail_ctx* ctx) {
    if (!ctx || !ctx->accum) return SAIL_ERR_INVALID;
    CU(cudaSetDevice(ctx->device));
    size_t n = npix(ctx);
    CU(cudaMemsetAsync(ctx->accum, 0, n * 3 * sizeof(float), ctx->stream));
    CU(cudaMemsetAsync(ctx->normal, 0, n * 3 * sizeof(float), ctx->stream));
    CU(cudaMemsetAsync(ctx->position, 0, n * 3 * sizeof(float), ctx->stream));
    CU(cudaMemsetAsync(ctx->depth, 0, n * sizeof(float), ctx->stream));
    CU(cudaMemsetAsync(ctx->hitId, 0xff, n * sizeof(int), ctx->stream));
    return SAIL_OK;
}

int sail_render_n(sail_ctx* ctx, const sail_frame* frames, int32_t count) {
    if (!ctx || !frames) return SAIL_ERR_INVALID;
    if (!ctx->sceneLoaded) return setErr(ctx, SAIL_ERR_STATE, "sail_render before sail_upload_scene");
    if (count <= 0 || count > SAIL_MAX_FRAMES_PER_CALL) return setErr(ctx, SAIL_ERR_INVALID, "frame count %d not in 1..%d", count, SAIL_MAX_FRAMES_PER_CALL);
    CU(cudaSetDevice(ctx->device));
    SailLaunch L;
    L.W = ctx->W; L.H = ctx->H; L.y0 = ctx->y0; L.y1 = ctx->y1;
    L.bandRows = ctx->bandRows; L.bandStride = ctx->bandStride; L.bandFirst = ctx->bandFirst;
    int tileRows = (ctx->y1 - ctx->y0 + SAIL_TILE_H - 1) / SAIL_TILE_H;
    if (ctx->bandStride > 1) {  // interleaved bands over the whole image
        int nbands = (ctx->H + ctx->bandRows - 1) / ctx->bandRows;
        int mine = (nbands - ctx->bandFirst + ctx->bandStride - 1) / ctx->bandStride;
        if (mine < 0) mine = 0;
        tileRows = mine * (ctx->bandRows / SAIL_TILE_H);
        L.y0 = 0; L.y1 = ctx->H;
        if (tileRows == 0) return SAIL_OK;
    }
    L.maxBounces = ctx->cfg.max_bounces;
    L.nframes = count;
    L.accumMode = ctx->cfg.accum_mode;
    L.writeAov = 1;
    L.rcpW = dm_rcp((float)ctx->W);
    L.rcpH = dm_rcp((float)ctx->H);
    L.accum = ctx->accum; L.normal = ctx->normal; L.position = ctx->position; L.hitId = ctx->hitId; L.depth = ctx->depth;
    L.scene = ctx->dScene;
    L.sceneSmem = (unsigned)sailSceneSmemBytes(ctx->hostScene.n, ctx->hostScene.ln);
    // timeSinceStart is seconds since the renderer started (tracer.js:98): never negative.  The cosine sampler relies on
    // seed = timeSinceStart + bounce >= 1 (sail_shade.cuh cosineSampleHemisphere).
    for (int k = 0; k < count; k++)
        if (!(frames[k].timeSinceStart >= 0.0f)) return setErr(ctx, SAIL_ERR_INVALID, "frame %d: timeSinceStart must be >= 0", k);
    for (int k = 0; k < count; k++) lowerFrame(frames[k], L.frames[k]);
    CU(cudaEventRecord(ctx->evStart, ctx->stream));
    // the variant follows the scene's content, never the caller's mask hints
    dim3 grid((ctx->W + SAIL_TILE_W - 1) / SAIL_TILE_W, tileRows);
    const int F = ctx->hostScene.features & SAIL_F_ALL;
    // SAIL_KERNEL_AUTO, from the measured matrix: the wavefront (512 paths per CTA when compiled at run time) pays
    // where the scene has many shape codes (quadrics); among the ahead-of-time variants also where it has many BSDFs (metal / glass).
    const bool tryJit = ctx->cfg.specialise != SAIL_SPECIALISE_OFF;
    bool wave = ctx->cfg.kernel == SAIL_KERNEL_WAVEFRONT;
    bool server = ctx->cfg.kernel == SAIL_KERNEL_SERVER;
    if (server && !tryJit) return setErr(ctx, SAIL_ERR_STATE, "SAIL_KERNEL_SERVER exists only as a scene-specialised kernel (specialise = off)");
    if (ctx->cfg.kernel == SAIL_KERNEL_AUTO) {
        wave = (F & SAIL_F_QUADRICS) != 0;
        // The server-warp kernel is never chosen automatically: measured slower than the plain megakernel on the scene it
        // was built for (Cornell box 9.3 vs 11.0 Gpaths/s); SAIL_B200_AUTO_SERVER=1 turns it on for A/B runs.
        static const bool autoServer = getenv("SAIL_B200_AUTO_SERVER") && getenv("SAIL_B200_AUTO_SERVER")[0] == '1';
        server = autoServer && tryJit && F == 0 && hasBoxesAndOthers(ctx->hostScene) && !(ctx->jitTried[2] && !ctx->jitFn[2]);
        if (server) wave = false;
        const bool jitFailed = tryJit && ctx->jitTried[wave ? 1 : 0] && !ctx->jitFn[wave ? 1 : 0];
        if (!server && (!tryJit || jitFailed)) wave = (F & (SAIL_F_QUADRICS | SAIL_F_GLOSSY)) != 0;
    }
    int kind = server ? 2 : (wave ? 1 : 0);
    void* jit = nullptr;
    if (tryJit) {
        for (int attempt = 0; attempt < 2 && !jit; attempt++) {
            if (!ctx->jitTried[kind]) {
                ctx->jitTried[kind] = true;
                SailJitKernel k;
                char why[300];
                why[0] = 0;
                if (sail_jit_get(ctx->device, kind, F, SAIL_MIN_CTAS(F), &ctx->hostScene, &k, why, sizeof(why)) == 0) {
                    ctx->jitFn[kind] = k.fn;
                    ctx->jitBlock[kind] = k.block;
                    ctx->jitPathWarps[kind] = k.pathWarps;
                    snprintf(ctx->jitNote[kind], sizeof(ctx->jitNote[kind]), "%s", k.info);
                } else {
                    ctx->jitFn[kind] = nullptr;
                    snprintf(ctx->jitNote[kind], sizeof(ctx->jitNote[kind]), "aot %s<F=%d> (jit: %s)", kind == 2 ? "mega" : (wave ? "wavefront" : "mega"), F, why);
                }
            }
            jit = ctx->jitFn[kind];
            if (!jit && kind == 2 && ctx->cfg.kernel == SAIL_KERNEL_AUTO) {  // no server kernel: the plain megakernel
                server = false;
                kind = 0;
                continue;
            }
            break;
        }
        if (!jit && kind == 2) return setErr(ctx, SAIL_ERR_STATE, "server kernel unavailable: %s", ctx->jitNote[2]);
        if (!jit && ctx->cfg.kernel == SAIL_KERNEL_AUTO && (F & SAIL_F_GLOSSY) && !wave) {  // first failure: re-decide for the AOT variants
            wave = true;
            kind = 1;
            snprintf(ctx->jitNote[1], sizeof(ctx->jitNote[1]), "aot wavefront<F=%d> (%s)", F, ctx->jitNote[0]);
            ctx->jitTried[1] = true;
        }
        if (!jit && ctx->cfg.specialise == SAIL_SPECIALISE_REQUIRE)
            return setErr(ctx, SAIL_ERR_STATE, "scene-specialised kernel unavailable: %s", ctx->jitNote[kind]);
        snprintf(ctx->kernelInfo, sizeof(ctx->kernelInfo), "%s", ctx->jitNote[kind]);
    } else {
        snprintf(ctx->kernelInfo, sizeof(ctx->kernelInfo), "aot %s<F=%d> (specialise = off)", wave ? "wavefront" : "mega", F);
    }
    if (jit) {
        char why[200];
        unsigned gx = grid.x, gy = grid.y;
        if (kind == 2) {  // strips of pathWarps 8x4 sub-tiles, one row group of 4 rows per grid.y
            const unsigned subW = (unsigned)(ctx->W + 7) / 8u, pw = (unsigned)ctx->jitPathWarps[2];
            gx = (subW + pw - 1) / pw;
            gy = (unsigned)tileRows * (SAIL_TILE_H / 4);
        }
        if (kind == 1) gx = ((unsigned)ctx->W + ctx->jitBlock[1] / 8u - 1u) / (ctx->jitBlock[1] / 8u);  // tile = (threads / 8) x 8
        if (sail_jit_launch(jit, &L, gx, gy, ctx->jitBlock[kind], L.sceneSmem, ctx->stream, why, sizeof(why)) != 0)
            return setErr(ctx, SAIL_ERR_CUDA, "%s", why);
    } else if (wave) CU(launchWave(F, L, grid, ctx->stream));
    else CU(launchMega(F, L, grid, ctx->stream));
    ctx->launches++;
    CU(cudaEventRecord(ctx->evStop, ctx->stream));
    ctx->timed = true;
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

int sail_synchronize(sail_ctx* ctx) {
    if (!ctx || !ctx->stream) return SAIL_ERR_INVALID;
    CU(cudaSetDevice(ctx->device));
    CU(cudaStreamSynchronize(ctx->stream));
    return SAIL_OK;
}

void* sail_device_ptr(sail_ctx* ctx, int32_t which) {
    if (!ctx) return nullptr;
    switch (which) {
        case SAIL_AOV_COLOR: return ctx->accum;
        case SAIL_AOV_NORMAL: return ctx->normal;
        case SAIL_AOV_POSITION: return ctx->position;
        case SAIL_AOV_HIT_ID: return ctx->hitId;
        case SAIL_AOV_DEPTH: return ctx->depth;
    }
    return nullptr;
}
void* sail_stream(sail_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }

int sail_read_aov(sail_ctx* ctx, int32_t which, void* out, size_t out_bytes) {
    if (!ctx || !out) return SAIL_ERR_INVALID;
    void* src = sail_device_ptr(ctx, which);
    if (!src) return setErr(ctx, SAIL_ERR_INVALID, "unknown AOV %d", which);
    size_t bytes = npix(ctx) * ((which == SAIL_AOV_HIT_ID || which == SAIL_AOV_DEPTH) ? 4 : 12);
    if (out_bytes != bytes) return setErr(ctx, SAIL_ERR_INVALID, "AOV %d needs %zu bytes, caller gave %zu", which, bytes, out_bytes);
    CU(cudaSetDevice(ctx->device));
    CU(cudaMemcpyAsync(out, src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return SAIL_OK;
}
int sail_read_accum(sail_ctx* ctx, float* rgb, size_t out_bytes) { return sail_read_aov(ctx, SAIL_AOV_COLOR, rgb, out_bytes); }

int sail_write_accum(sail_ctx* ctx, const float* rgb, size_t in_bytes) {
    if (!ctx || !rgb) return SAIL_ERR_INVALID;
    size_t bytes = npix(ctx) * 12;
    if (in_bytes != bytes) return setErr(ctx, SAIL_ERR_INVALID, "accumulation buffer is %zu bytes, caller gave %zu", bytes, in_bytes);
    CU(cudaSetDevice(ctx->device));
    CU(cudaMemcpyAsync(ctx->accum, rgb, bytes, cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return SAIL_OK;
}

int sail_scale_accum(sail_ctx* ctx, float s) {
    if (!ctx || !ctx->accum) return SAIL_ERR_INVALID;
    CU(cudaSetDevice(ctx->device));
    size_t n = npix(ctx) * 3;
    sail_scale_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(ctx->accum, n, s);
    ctx->launches++;
    CU(cudaGetLastError());
    return SAIL_OK;
}

namespace {
void fillPresentParams(sail_ctx* ctx, PresentParams& P) {
    P.W = ctx->W; P.H = ctx->H;
    P.filter_id = ctx->cfg.filter_id;
    P.gamma_c = ctx->cfg.filter_gamma_c;
    P.radius[0] = ctx->cfg.filter_radius[0]; P.radius[1] = ctx->cfg.filter_radius[1];
    memcpy(P.weights, ctx->cfg.filter_weights, sizeof(P.weights));
    P.color = ctx->accum; P.normal = ctx->normal; P.position = ctx->position;
    P.out_f32 = nullptr;
    P.out_u8 = nullptr;
}
}  // namespace

int sail_present_begin(sail_ctx* ctx, uint8_t* rgba8) {
    if (!ctx || !ctx->accum || !rgba8) return SAIL_ERR_INVALID;
    if (ctx->copyInFlight) return setErr(ctx, SAIL_ERR_STATE, "sail_present_begin: the previous image has not been collected with sail_present_end");
    CU(cudaSetDevice(ctx->device));
    size_t n = npix(ctx);
    if (!ctx->presentU8) CU(cudaMalloc(&ctx->presentU8, n * 4));
    PresentParams P;
    fillPresentParams(ctx, P);
    P.out_u8 = ctx->presentU8;
    // the filter overwrites the device image: it must not overtake the previous copy (long finished in practice)
    if (ctx->copiedOnce) CU(cudaStreamWaitEvent(ctx->stream, ctx->evCopied, 0));
    dim3 grid((ctx->W + 31) / 32, (ctx->H + 7) / 8);
    sail_present_kernel<<<grid, 256, 0, ctx->stream>>>(P);
    ctx->launches++;
    CU(cudaGetLastError());
    CU(cudaEventRecord(ctx->evPresented, ctx->stream));
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
    CU(cudaSetDevice(ctx->device));
    CU(cudaEventSynchronize(ctx->evCopied));
    ctx->copyInFlight = false;
    return SAIL_OK;
}

int sail_present(sail_ctx* ctx, float* rgba_f32, uint8_t* rgba8) {
    if (!ctx || !ctx->accum) return SAIL_ERR_INVALID;
    CU(cudaSetDevice(ctx->device));
    size_t n = npix(ctx);
    if (rgba_f32 && !ctx->presentF32) CU(cudaMalloc(&ctx->presentF32, n * 4 * sizeof(float)));
    if (rgba8 && !ctx->presentU8) CU(cudaMalloc(&ctx->presentU8, n * 4));
    if (rgba8 && ctx->copyInFlight) return setErr(ctx, SAIL_ERR_STATE, "sail_present while a sail_present_begin image is in flight");
    PresentParams P;
    P.W = ctx->W; P.H = ctx->H;
    P.filter_id = ctx->cfg.filter_id;
    P.gamma_c = ctx->cfg.filter_gamma_c;
    P.radius[0] = ctx->cfg.filter_radius[0]; P.radius[1] = ctx->cfg.filter_radius[1];
    memcpy(P.weights, ctx->cfg.filter_weights, sizeof(P.weights));
    P.color = ctx->accum; P.normal = ctx->normal; P.position = ctx->position;
    P.out_f32 = rgba_f32 ? ctx->presentF32 : nullptr;
    P.out_u8 = rgba8 ? ctx->presentU8 : nullptr;
    dim3 grid((ctx->W + 31) / 32, (ctx->H + 7) / 8);
    sail_present_kernel<<<grid, 256, 0, ctx->stream>>>(P);
    ctx->launches++;
    CU(cudaGetLastError());
    if (rgba_f32) CU(cudaMemcpyAsync(rgba_f32, ctx->presentF32, n * 4 * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    if (rgba8) CU(cudaMemcpyAsync(rgba8, ctx->presentU8, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return SAIL_OK;
}

int sail_intersect_rays(sail_ctx* ctx, const float* rays, int32_t count, int32_t* hit_id, float* t, float* normal, int32_t* mat_category) {
    if (!ctx || !rays || count < 0) return SAIL_ERR_INVALID;
    if (!ctx->sceneLoaded) return setErr(ctx, SAIL_ERR_STATE, "sail_intersect_rays before sail_upload_scene");
    if (count == 0) return SAIL_OK;
    CU(cudaSetDevice(ctx->device));
    float *dR = nullptr, *dT = nullptr, *dN = nullptr;
    int *dI = nullptr, *dM = nullptr;
    size_t c = (size_t)count;
    CU(cudaMalloc(&dR, c * 6 * sizeof(float)));
    CU(cudaMalloc(&dT, c * sizeof(float)));
    CU(cudaMalloc(&dN, c * 3 * sizeof(float)));
    CU(cudaMalloc(&dI, c * sizeof(int)));
    CU(cudaMalloc(&dM, c * sizeof(int)));
    CU(cudaMemcpyAsync(dR, rays, c * 6 * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    sail_intersect_kernel<<<(unsigned)((c + SAIL_CTA_THREADS - 1) / SAIL_CTA_THREADS), SAIL_CTA_THREADS, sailSceneSmemBytes(ctx->hostScene.n, ctx->hostScene.ln), ctx->stream>>>(ctx->dScene, dR, count, dI, dT, dN, dM);
    ctx->launches++;
    CU(cudaGetLastError());
    if (hit_id) CU(cudaMemcpyAsync(hit_id, dI, c * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    if (t) CU(cudaMemcpyAsync(t, dT, c * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    if (normal) CU(cudaMemcpyAsync(normal, dN, c * 3 * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    if (mat_category) CU(cudaMemcpyAsync(mat_category, dM, c * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    cudaFree(dR); cudaFree(dT); cudaFree(dN); cudaFree(dI); cudaFree(dM);
    return SAIL_OK;
}

int sail_trace_dump(sail_ctx* ctx, const sail_frame* frame, int32_t* hit_id, float* t, float* rays, float* radiance) {
    if (!ctx || !frame) return SAIL_ERR_INVALID;
    if (!ctx->sceneLoaded) return setErr(ctx, SAIL_ERR_STATE, "sail_trace_dump before sail_upload_scene");
    CU(cudaSetDevice(ctx->device));
    size_t n = npix(ctx), B = (size_t)ctx->cfg.max_bounces;
    if (B == 0) return setErr(ctx, SAIL_ERR_INVALID, "max_bounces is 0");
    SailDumpLaunch L;
    L.W = ctx->W; L.H = ctx->H; L.maxBounces = (int)B;
    L.rcpW = dm_rcp((float)ctx->W); L.rcpH = dm_rcp((float)ctx->H);
    L.scene = ctx->dScene;
    L.hitId = nullptr; L.t = nullptr; L.rays = nullptr; L.radiance = nullptr;
    if (!(frame->timeSinceStart >= 0.0f)) return setErr(ctx, SAIL_ERR_INVALID, "timeSinceStart must be >= 0");
    lowerFrame(*frame, L.frame);
    if (hit_id) CU(cudaMalloc(&L.hitId, n * B * sizeof(int)));
    if (t) CU(cudaMalloc(&L.t, n * B * sizeof(float)));
    if (rays) CU(cudaMalloc(&L.rays, n * B * 6 * sizeof(float)));
    if (radiance) CU(cudaMalloc(&L.radiance, n * 3 * sizeof(float)));
    dim3 grid((ctx->W + SAIL_TILE_W - 1) / SAIL_TILE_W, (ctx->H + SAIL_TILE_H - 1) / SAIL_TILE_H);
    sail_trace_dump_kernel<<<grid, SAIL_CTA_THREADS, sailSceneSmemBytes(ctx->hostScene.n, ctx->hostScene.ln), ctx->stream>>>(L);
    ctx->launches++;
    CU(cudaGetLastError());
    if (hit_id) CU(cudaMemcpyAsync(hit_id, L.hitId, n * B * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    if (t) CU(cudaMemcpyAsync(t, L.t, n * B * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    if (rays) CU(cudaMemcpyAsync(rays, L.rays, n * B * 6 * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    if (radiance) CU(cudaMemcpyAsync(radiance, L.radiance, n * 3 * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    cudaFree(L.hitId); cudaFree(L.t); cudaFree(L.rays); cudaFree(L.radiance);
    return SAIL_OK;
}

int sail_eval_math(sail_ctx* ctx, int32_t which, const float* a, const float* b, float* out, int32_t count) {
    if (!ctx || !a || !out || count < 0 || which < 0 || which > 12) return SAIL_ERR_INVALID;
    if (count == 0) return SAIL_OK;
    CU(cudaSetDevice(ctx->device));
    float *dA = nullptr, *dB = nullptr, *dO = nullptr;
    size_t bytes = (size_t)count * sizeof(float);
    CU(cudaMalloc(&dA, bytes));
    CU(cudaMalloc(&dO, bytes));
    CU(cudaMemcpyAsync(dA, a, bytes, cudaMemcpyHostToDevice, ctx->stream));
    if (b) {
        CU(cudaMalloc(&dB, bytes));
        CU(cudaMemcpyAsync(dB, b, bytes, cudaMemcpyHostToDevice, ctx->stream));
    }
    sail_eval_math_kernel<<<(count + 255) / 256, 256, 0, ctx->stream>>>(which, dA, dB, dO, count);
    ctx->launches++;
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(out, dO, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    cudaFree(dA); cudaFree(dB); cudaFree(dO);
    return SAIL_OK;
}

int sail_selftest_math(sail_ctx* ctx, uint64_t* mismatches) {
    if (!ctx || !mismatches) return SAIL_ERR_INVALID;
    CU(cudaSetDevice(ctx->device));
    unsigned long long* d = nullptr;
    CU(cudaMalloc(&d, sizeof(unsigned long long)));
    CU(cudaMemsetAsync(d, 0, sizeof(unsigned long long), ctx->stream));
    sail_selftest_rcp_kernel<<<(1u << 20) / 256, 256, 0, ctx->stream>>>(d);
    sail_selftest_unit_kernel<<<(1u << 19) / 256, 256, 0, ctx->stream>>>(d);
    ctx->launches++;
    CU(cudaGetLastError());
    unsigned long long h = 0;
    CU(cudaMemcpyAsync(&h, d, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    cudaFree(d);
    *mismatches = h;
    return SAIL_OK;
}

uint64_t sail_launch_count(const sail_ctx* ctx) { return ctx ? ctx->launches : 0; }

int sail_last_render_ms(sail_ctx* ctx, float* ms) {
    if (!ctx || !ms) return SAIL_ERR_INVALID;
    if (!ctx->timed) return setErr(ctx, SAIL_ERR_STATE, "no render has been issued yet");
    CU(cudaSetDevice(ctx->device));
    CU(cudaEventSynchronize(ctx->evStop));
    CU(cudaEventElapsedTime(ms, ctx->evStart, ctx->evStop));
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
