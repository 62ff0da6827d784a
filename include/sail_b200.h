/* sail_b200.h -- C ABI of libsail_b200.so: the B200 (sm_100a) path-tracing core behind Sail's
 * Renderer / Tracer.
 *
 * This is the drop-in boundary (DESIGN.md "Boundary").  Each entry point replaces one
 * step of the reference's WebGL path; the citation is the reference code that the function stands in
 * for (paths relative to /root/reference).  The host (Node.js through the N-API shim of
 * INTEGRATION.md, or the Python mirror in sail_b200/) keeps building the same three Float32 tables and
 * the same per-frame uniforms as src/core/tracer.js does and hands them over unchanged.
 *
 * Conventions: plain C types only; every function returns 0 on success or a negative SAIL_ERR_*;
 * sail_last_error() gives the message of the last failure on that context.  A context is not thread
 * safe (the reference is driven from one JS event loop).  sail_render*() are asynchronous on the
 * context's CUDA stream; sail_read_*() / sail_present() synchronise.  Image row 0 is the BOTTOM row
 * (gl_FragCoord convention); pixels are row-major, interleaved channels.  There is no CPU fallback:
 * without a usable sm_100 device sail_create() fails with SAIL_ERR_CUDA.
 */
#ifndef SAIL_B200_H
#define SAIL_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SAIL_ABI_VERSION 2

/* Row widths of the wire format (src/core/webgl.js:137-139). */
#define SAIL_OBJECTS_LENGTH 18
#define SAIL_TEXPARAMS_LENGTH 16
#define SAIL_LIGHTS_LENGTH 18
#define SAIL_MAX_FRAMES_PER_CALL 32
#define SAIL_MAX_DEVICES 8
#define SAIL_GROUP_ID_BYTES 128

enum {
    SAIL_OK = 0,
    SAIL_ERR_CUDA = -1,     /* no device / CUDA runtime failure (reference: alert(), src/core/webgl.js:222-225) */
    SAIL_ERR_INVALID = -2,  /* bad argument or malformed table (reference: throw "<string>", src/scene/light.js:52) */
    SAIL_ERR_STATE = -3,    /* call out of order, e.g. render before upload (reference: silent no-op, webgl.js:52) */
    SAIL_ERR_NOMEM = -4
};

/* Display filter plugins (src/shader/filter/shader.filter.js:18-30). */
enum {
    SAIL_FILTER_COLOR = 0, SAIL_FILTER_GAMMA = 1, SAIL_FILTER_BOX = 2, SAIL_FILTER_GAUSSIAN = 3,
    SAIL_FILTER_MITCHELL = 4, SAIL_FILTER_SINC = 5, SAIL_FILTER_TRIANGLE = 6, SAIL_FILTER_NORMAL = 7,
    SAIL_FILTER_POSITION = 8, SAIL_FILTER_WAVELET = 9, SAIL_FILTER_TONEMAPPING = 10
};

/* Per-pixel outputs.  COLOR/NORMAL/POSITION are the three MRT attachments of
 * src/shader/main/fstrace.glsl:2-4,14-16; HIT_ID and DEPTH expose Intersect.index / Intersect.d of the
 * primary ray (src/shader/const/struct.glsl:2,11), which the reference computes but never outputs. */
enum {
    SAIL_AOV_COLOR = 0,     /* float32 x3: the progressive accumulation (frameCache[0], src/core/webgl.js:88-92) */
    SAIL_AOV_NORMAL = 1,    /* float32 x3: n/2 + 0.5 */
    SAIL_AOV_POSITION = 2,  /* float32 x3: normalize(p) */
    SAIL_AOV_HIT_ID = 3,    /* int32 x1: object index of the first hit, -1 = miss */
    SAIL_AOV_DEPTH = 4      /* float32 x1: first-hit distance in units of the un-normalised primary ray */
};

enum {
    SAIL_ACCUM_MEAN = 0,  /* out = mix(e, cache, textureWeight)  (fstrace.glsl:13-14) -- the reference's update */
    SAIL_ACCUM_SUM = 1    /* out = cache + e; textureWeight ignored.  For sample-sharded multi-GPU rendering:
                             per-GPU sums are reduced and divided by the sample count (DESIGN.md "Multi-GPU") */
};

enum {
    SAIL_KERNEL_MEGA = 0,       /* one thread traces a whole path */
    SAIL_KERNEL_WAVEFRONT = 1,  /* per-bounce stages over ray queues regrouped by hit object, queues in shared memory */
    SAIL_KERNEL_AUTO = 2,       /* default: chosen from the scene's content by the measured policy of DESIGN.md 4.2 */
    SAIL_KERNEL_QUEUE = 3       /* wavefront with its ray queues in device memory: raygen / scan / sort / shade / accumulate as separate
                                   kernels over batches of up to 4 M paths regrouped globally by hit object (sail_queue.cuh).  Exists
                                   as a scene-specialised build only, for contexts that render whole images; otherwise the
                                   shared-memory wavefront runs (sail_kernel_info() says so). */
};

enum {
    SAIL_SPECIALISE_AUTO = 0,    /* default: compile the kernel for the scene's structure (NVRTC; object count and per-slot shape
                                    categories become constants -- the per-scene shader generation of
                                    src/shader/generator.js:107-123) on a worker thread; frames rendered before it is ready use the
                                    resident ahead-of-time variant (bit-identical), as they do when compiling is not possible.  The
                                    reference blocks in createProgram instead (src/core/webgl.js:165-213). */
    SAIL_SPECIALISE_OFF = 1,     /* ahead-of-time variants only */
    SAIL_SPECIALISE_REQUIRE = 2  /* compile synchronously at the first sail_render* after an upload and fail with SAIL_ERR_STATE if the
                                    scene-specialised kernel cannot be built */
};

/* How the images of a multi-GPU context are split (the reference has one GL context, src/core/webgl.js:215-227). */
enum {
    SAIL_SHARD_NONE = 0,
    SAIL_SHARD_SAMPLE = 1,  /* member g traces the frames whose index (counted since the last restart) is congruent to g modulo the
                               group size, into its own FP32 sum; the display / read-back calls combine the sums (BASELINE config 4) */
    SAIL_SHARD_TILE = 2     /* member g traces every frame for its interleaved row bands; display / read-back gather the rows (config 5) */
};
/* How a sharded context combines its members' buffers when an image is asked for. */
enum {
    SAIL_COMBINE_AUTO = 0,  /* peer memory when every member can map the root's and its peers' buffers, NCCL otherwise */
    SAIL_COMBINE_NCCL = 1,  /* ncclReduce(sum) to the root (libnccl.so.2 is opened with dlopen; no link-time dependency) */
    SAIL_COMBINE_P2P = 2    /* members add / copy their share straight into the root's buffer over NVLink peer mappings, in a fixed
                               member order (deterministic sums) */
};

typedef struct sail_ctx sail_ctx;

/* new Sail.Renderer(canvas): src/core/renderer.js:9-39 + WebglHelper.initWebgl (src/core/webgl.js:215-227)
 * + the frameCache allocation (webgl.js:12-28).  The 512x512 canvas becomes an explicit size. */
typedef struct sail_create_info {
    int32_t width, height;  /* full image size in pixels */
    int32_t device;         /* CUDA device ordinal (ndev <= 1) */
    int32_t row_begin;      /* rows [row_begin, row_end) are rendered by this context (host-managed tile sharding); */
    int32_t row_end;        /* 0,0 = the whole image.  Buffers always cover the full image. */
    /* ABI 2: one context over several GPUs of this process.  The context owns a stream, the buffers and the communicator of every
     * member; device_ids[0] is the root, where combined images are formed, filtered and read back. */
    int32_t ndev;           /* 0 or 1: single device.  2..SAIL_MAX_DEVICES: device_ids[0..ndev) */
    int32_t shard_mode;     /* SAIL_SHARD_SAMPLE or SAIL_SHARD_TILE when ndev > 1 */
    int32_t band_rows;      /* SAIL_SHARD_TILE: rows per interleaved band, a multiple of 8 (0 = 8, the measured best: profiles/r02f_scaling.json) */
    int32_t device_ids[SAIL_MAX_DEVICES];  /* a device may appear more than once: those members share the GPU (peer-memory combine only) */
    int32_t combine;        /* SAIL_COMBINE_* */
    int32_t reserved[3];
} sail_create_info;

/* scene.tracerConfig() / scene.rendererConfig() lowered to PODs (src/scene/scene.js:70-118; plugin
 * parameters of src/shader/generator.js:10-24).  The masks are specialisation hints (bit = category id of
 * src/shader/const/define.glsl:18-44); results never depend on them. */
typedef struct sail_config {
    uint32_t shape_mask, material_mask, texture_mask, light_mask;
    int32_t trace_id;         /* 0 = "path" (src/shader/trace/shader.trace.js:7-9) */
    int32_t max_bounces;      /* MAXBOUNCES, const/define.glsl:6 (5) */
    int32_t accum_mode;       /* SAIL_ACCUM_* */
    int32_t kernel;           /* SAIL_KERNEL_* */
    int32_t filter_id;        /* SAIL_FILTER_* */
    float filter_gamma_c;     /* FILTER_GAMMA_C (filter/gamma.glsl:1-8) */
    float filter_radius[2];   /* FILTER_WINDOW_RADIUS / FILTER_WAVELET_R */
    float filter_weights[16]; /* windowWeightTable built by filter/{box,gaussian,mitchell,sinc,triangle}.js */
    int32_t specialise;       /* SAIL_SPECIALISE_* (was reserved: 0 keeps the default) */
    int32_t rr_start_bounce;  /* Russian roulette from this bounce on: a path survives bounce d >= rr_start_bounce with probability
                                 q = clamp(max(throughput), 0.05, 1) and is scaled by 1/q.  0 = off (default).  NOT in the reference
                                 (fixed MAXBOUNCES, const/define.glsl:6): an opt-in extension; images then differ from the
                                 reference estimator's by noise, not in expectation. */
    int32_t reserved[2];
} sail_config;

/* One frame's uniforms: Tracer.render(), src/core/tracer.js:92-101 / src/core/shader.js:41-49. */
typedef struct sail_frame {
    float matrix[16];     /* uniform mat4 matrix, column-major, already inverted and jittered on the host */
    float eye[3];         /* uniform vec3 eye */
    float textureWeight;  /* sampleCount / (sampleCount + 1) */
    float timeSinceStart; /* seconds since the renderer started, >= 0 (SAIL_ERR_INVALID otherwise); seeds the per-pixel hash RNG (src/shader/trace/path.glsl:21) */
} sail_frame;

const char* sail_version(void);
int sail_abi_version(void);
/* Number of usable sm_100 devices (0 = none; never falls back to the CPU). */
int sail_device_count(void);

int sail_create(const sail_create_info* info, sail_ctx** out);
void sail_destroy(sail_ctx* ctx);

/* One GPU per process (torchrun-style hosts): every process creates a single-device context and joins the same group; the
 * context then is member `rank` of `nranks` and behaves like one device of a sharded context.  sail_group_id() makes the
 * 128-byte rendezvous token (ncclGetUniqueId) on one process; the host carries it to the others (any channel: a file, MPI,
 * torch.distributed ...).  After the join every process issues the SAME sequence of sail_upload_* / sail_render* /
 * sail_present* / sail_read_* / sail_reset calls with the same arguments (they are collective); images arrive on rank 0,
 * the other ranks' output pointers may be NULL.  sail_group_join is collective and blocks until all ranks have joined. */
int sail_group_id(uint8_t id[SAIL_GROUP_ID_BYTES]);
int sail_group_join(sail_ctx* ctx, int32_t nranks, int32_t rank, const uint8_t id[SAIL_GROUP_ID_BYTES], int32_t shard_mode,
                    int32_t band_rows, int32_t combine);
/* Group size, this context's first member index, and a description of the combine path ("p2p", "nccl 2.28.9", ...). */
int sail_group_info(sail_ctx* ctx, int32_t* nmembers, int32_t* first_member, char* buf, int32_t len);
/* Device time of the most recent combine (reduce / gather to the root) in milliseconds, on the root's stream. */
int sail_last_combine_ms(sail_ctx* ctx, float* ms);
const char* sail_last_error(const sail_ctx* ctx);
void sail_default_config(sail_config* cfg);

/* setProgram(new TraceShader(scene.tracerConfig())) + setProgram(new RenderShader(scene.rendererConfig())):
 * src/core/tracer.js:43, src/core/renderer.js:46. */
int sail_configure(sail_ctx* ctx, const sail_config* cfg);

/* Tracer.update(scene): src/core/tracer.js:42-90.  The library copies the tables (as texImage2D does). */
int sail_upload_scene(sail_ctx* ctx, const float* objects, int32_t n, const float* texparams, int32_t tn,
                      const float* lights, int32_t ln);
/* Tracer.updateObjects(scene): src/core/tracer.js:25-40 (objects table only; n may change). */
int sail_upload_objects(sail_ctx* ctx, const float* objects, int32_t n);

/* Validates the three tables exactly as sail_upload_scene() would (no device needed) and returns the
 * kernel feature bits the scene content selects (1 quadrics, 2 metal/glass, 4 uv textures, 8 lights), or a
 * negative SAIL_ERR_*.  `err` (may be NULL) receives the message. */
int sail_scene_features(const float* objects, int32_t n, const float* texparams, int32_t tn, const float* lights,
                        int32_t ln, char* err, int32_t errlen);

/* Tile sharding with interleaved bands: this context renders the rows y with
 * (y / band_rows) % band_stride == band_first (band_rows a multiple of 8).  band_stride = 1 restores the
 * contiguous [row_begin, row_end) range of sail_create(). */
int sail_set_row_bands(sail_ctx* ctx, int32_t band_rows, int32_t band_stride, int32_t band_first);

/* Clears the accumulation and AOV buffers (a fresh frameCache; the reference instead restarts the running
 * mean with textureWeight = 0 when scene.sampleCount is reset, src/core/renderer.js:57-60). */
int sail_reset(sail_ctx* ctx);

/* Tracer.render(mvp, eye, sampleCount) -> shader.render('triangle'): src/core/tracer.js:92-101,
 * src/core/webgl.js:51-93.  One call = 1 sample per pixel over this context's rows. */
int sail_render(sail_ctx* ctx, const float matrix_colmajor[16], const float eye[3], float textureWeight,
                float timeSinceStart);
/* `count` consecutive frames (1..SAIL_MAX_FRAMES_PER_CALL; a sample-sharded context takes that many per member, i.e. up to
 * SAIL_MAX_FRAMES_PER_CALL x group size); pixel results equal `count` sail_render() calls, the accumulator stays in
 * registers between frames.  AOVs are those of the last frame, as they would be after `count` calls.
 * Sharded contexts: collective (every member of a joined group passes the same frames and traces its share); with
 * SAIL_SHARD_SAMPLE and SAIL_ACCUM_MEAN a frame with textureWeight == 0 restarts the sums, as it restarts the running mean. */
int sail_render_n(sail_ctx* ctx, const sail_frame* frames, int32_t count);

int sail_synchronize(sail_ctx* ctx);
/* Blocks until the scene-specialised kernel requested by the last upload / render has been compiled (or has failed) and, when
 * it is ready, makes the next sail_render* use it.  Returns 0 when that kernel is in place, 1 when the context keeps running
 * the ahead-of-time variant (sail_kernel_info() says why).  Benchmarks call it once after the first frame. */
int sail_specialise_wait(sail_ctx* ctx);

/* Read back a full-image buffer (what readPixels on frameCache[0], [2], [3] would give). */
int sail_read_aov(sail_ctx* ctx, int32_t which, void* out, size_t out_bytes);
int sail_read_accum(sail_ctx* ctx, float* rgb, size_t out_bytes);
/* Overwrite the accumulation buffer (resume a saved progressive image / install a reduced multi-GPU sum). */
int sail_write_accum(sail_ctx* ctx, const float* rgb, size_t in_bytes);
/* accum *= s (SAIL_ACCUM_SUM: divide the reduced sum by the sample count). */
int sail_scale_accum(sail_ctx* ctx, float s);

/* renderShader.render('triangle'): src/core/renderer.js:63, src/shader/main/fsrender.glsl:4-6 with the
 * configured filter plugin.  Either output may be NULL.  rgba_f32: W*H*4 floats; rgba8: W*H*4 bytes
 * (what the canvas would hold: clamped to [0,1], rounded to 8 bits). */
int sail_present(sail_ctx* ctx, float* rgba_f32, uint8_t* rgba8);

/* The same display pass as a swapchain of two, for hosts that keep rendering while an image travels to the host -- what the
 * browser's compositor does with the canvas of src/core/renderer.js:63 between two requestAnimationFrame callbacks
 * (ui/ui.js:67-70).  sail_present_begin() queues the display filter behind the frames rendered so far and starts the
 * device-to-host copy of the RGBA8 image on a second stream; it does not wait, and the next sail_render*() may be issued
 * at once (it overlaps the copy).  sail_present_end() waits until `rgba8` (page-locked memory from sail_alloc_host for a
 * truly asynchronous copy) holds the image.  One image in flight per context: begin, end, begin, end ... */
int sail_present_begin(sail_ctx* ctx, uint8_t* rgba8);
int sail_present_end(sail_ctx* ctx);

/* Device-side handles for hosts that combine buffers across GPUs themselves (torch.distributed / NCCL on
 * the host side): raw device pointers of the full-image buffers and the CUDA stream (cudaStream_t).  The normal / position /
 * hit-id / depth buffers are filled on demand (by this call, sail_read_aov and the filters that sample them), on that stream. */
void* sail_device_ptr(sail_ctx* ctx, int32_t which);
void* sail_stream(sail_ctx* ctx);

/* Diagnostics / picking: intersectObjects() (src/shader/shape/shader.shape.js:28-51) on caller-supplied
 * rays.  rays: count x 6 (origin, dir).  Outputs (each may be NULL): hit_id (-1 = miss), t,
 * normal count x 3 (flipped toward the ray), mat_category. */
int sail_intersect_rays(sail_ctx* ctx, const float* rays, int32_t count, int32_t* hit_id, float* t, float* normal,
                        int32_t* mat_category);
/* sail_intersect_rays() for callers that must not allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same
 * arguments, same results; the device copies of the rays and outputs live in a fixed array inside the library instead of
 * per-call cudaMalloc / cudaFree, so count <= SAIL_NOALLOC_MAX_RAYS (more: SAIL_ERR_INVALID, nothing is launched).  The call
 * makes no cudaMalloc / cudaFree and calls no allocating library routine.  The array (3 MiB per GPU that loads the library,
 * held until exit) is shared by every call on a device and keeps the last call's contents: calls that overlap on the GPU --
 * two contexts on one device driven from two threads -- race on it; calls on one context are ordered by its stream. */
#define SAIL_NOALLOC_MAX_RAYS 65536
int sail_intersect_rays_noalloc(sail_ctx* ctx, const float* rays, int32_t count, int32_t* hit_id, float* t, float* normal,
                                int32_t* mat_category);
/* One frame with per-bounce records (no accumulation): hit_id/t are W*H*max_bounces (unreached = -2),
 * rays W*H*max_bounces*6, radiance W*H*3 (the raw sample e).  Each may be NULL. */
int sail_trace_dump(sail_ctx* ctx, const sail_frame* frame, int32_t* hit_id, float* t, float* rays, float* radiance);
/* Elementwise evaluation of the device math library: which = 0 sin, 1 cos, 2 tan, 3 atan, 4 acos, 5 asin,
 * 6 exp, 7 log, 8 atan(a,b), 9 pow(a,b), 10 a/b, 11 sqrt, 12 1/a. */
int sail_eval_math(sail_ctx* ctx, int32_t which, const float* a, const float* b, float* out, int32_t count);
/* Runs the device reciprocal over all 2^32 float inputs and counts results that differ from its stated rule
 * (correctly rounded inside 2^-126 <= |a| <= 2^126, pinned specials outside; sail_b200/csrc/detmath.h). */
int sail_selftest_math(sail_ctx* ctx, uint64_t* mismatches);
/* Number of kernels this context has launched since creation. */
uint64_t sail_launch_count(const sail_ctx* ctx);
/* Device time of the most recent sail_render() / sail_render_n() call, from CUDA events recorded on the
 * context's stream around its kernels, in milliseconds (synchronises). */
int sail_last_render_ms(sail_ctx* ctx, float* ms);
/* One line describing the trace kernel the most recent sail_render* used, e.g.
 * "jit mega<F=0> n=3 cats=1,9,2 (nvrtc 12.9, 0.8 s)" or "aot wavefront<F=10> (jit: NVRTC not found ...)".
 * Returns the length written (excluding the terminator), or a negative error code. */
int sail_kernel_info(sail_ctx* ctx, char* buf, int32_t len);
/* Compiles the scene-specialised kernel for these tables without loading it: needs NVRTC but neither a GPU nor the
 * driver library, so a build machine can check that a scene's kernel compiles.  kernel = SAIL_KERNEL_MEGA / _WAVEFRONT.
 * Returns 0 and a description in `info`, or a negative code and the reason (NVRTC missing, compiler log, bad tables). */
int sail_jit_compile_check(const float* objects, int32_t n, const float* texparams, int32_t tn, const float* lights, int32_t ln,
                           int32_t kernel, char* info, int32_t infolen);
/* Page-locked host memory for the caller's image / table buffers, so that sail_present / sail_read_* copy
 * straight into it at full PCIe rate (the destination ArrayBuffer of the reference's gl.readPixels,
 * src/core/webgl.js:88-92 usage in renderer.js; an N-API host wraps it as an external ArrayBuffer).  Returns NULL
 * on failure.  Buffers from any other allocator are accepted everywhere too (staged by the driver, slower). */
void* sail_alloc_host(uint64_t bytes);
void sail_free_host(void* p);

#ifdef __cplusplus
}
#endif
#endif
