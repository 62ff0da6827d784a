// sail_napi.cc -- N-API addon over the C ABI of include/sail_b200.h.
// Build with node-gyp (binding.gyp); NOT compiled on the build image (no node_api.h there).
// Each function unwraps the context handle and the typed arrays, calls the C ABI and turns a non-zero return
// into a JavaScript Error carrying sail_last_error().
#include <node_api.h>
#include <stdint.h>
#include <string.h>

#include "sail_b200.h"

#define NAPI_OK(call) do { if ((call) != napi_ok) { napi_throw_error(env, nullptr, #call " failed"); return nullptr; } } while (0)

static bool check(napi_env env, sail_ctx* ctx, int rc) {
    if (rc == SAIL_OK) return true;
    napi_throw_error(env, nullptr, ctx ? sail_last_error(ctx) : "sail_b200: invalid call");
    return false;
}
static sail_ctx* handle(napi_env env, napi_value v) {
    void* p = nullptr;
    napi_get_value_external(env, v, &p);
    return (sail_ctx*)p;
}
static float* f32(napi_env env, napi_value v, size_t* len) {
    napi_typedarray_type type; napi_value buf; size_t off; void* data = nullptr;
    if (napi_get_typedarray_info(env, v, &type, len, &data, &buf, &off) != napi_ok || type != napi_float32_array) {
        napi_throw_type_error(env, nullptr, "expected a Float32Array");
        *len = 0;
        return nullptr;
    }
    return (float*)data;
}
static double numberProp(napi_env env, napi_value obj, const char* name) {
    napi_value v; double d = 0;
    napi_get_named_property(env, obj, name, &v);
    napi_get_value_double(env, v, &d);
    return d;
}
static void floatArrayProp(napi_env env, napi_value obj, const char* name, float* out, uint32_t n) {
    napi_value arr, e;
    napi_get_named_property(env, obj, name, &arr);
    for (uint32_t i = 0; i < n; i++) { double d = 0; napi_get_element(env, arr, i, &e); napi_get_value_double(env, e, &d); out[i] = (float)d; }
}

// create(width, height, device | {devices: [0, 1, ...], shard: 1 | 2, bandRows, combine}) -> {handle, width, height}
// [new Sail.Renderer(canvas)].  With a device list the context is sharded over those GPUs behind the C ABI
// (sail_create_info.ndev / device_ids / shard_mode): sample index mod N (shard 1) or interleaved row bands (shard 2).
static napi_value Create(napi_env env, napi_callback_info info) {
    size_t argc = 3; napi_value a[3];
    NAPI_OK(napi_get_cb_info(env, info, &argc, a, nullptr, nullptr));
    sail_create_info ci; memset(&ci, 0, sizeof(ci));
    napi_get_value_int32(env, a[0], &ci.width);
    napi_get_value_int32(env, a[1], &ci.height);
    if (argc > 2) {
        napi_valuetype vt;
        napi_typeof(env, a[2], &vt);
        if (vt == napi_object) {
            napi_value devs; uint32_t n = 0;
            napi_get_named_property(env, a[2], "devices", &devs);
            napi_get_array_length(env, devs, &n);
            if (n > SAIL_MAX_DEVICES) { napi_throw_range_error(env, nullptr, "too many devices"); return nullptr; }
            ci.ndev = (int32_t)n;
            for (uint32_t i = 0; i < n; i++) { napi_value e; napi_get_element(env, devs, i, &e); napi_get_value_int32(env, e, &ci.device_ids[i]); }
            ci.shard_mode = (int32_t)numberProp(env, a[2], "shard");
            ci.band_rows = (int32_t)numberProp(env, a[2], "bandRows");
            ci.combine = (int32_t)numberProp(env, a[2], "combine");
        } else {
            napi_get_value_int32(env, a[2], &ci.device);
        }
    }
    sail_ctx* ctx = nullptr;
    int rc = sail_create(&ci, &ctx);
    if (rc != SAIL_OK) { napi_throw_error(env, nullptr, ctx ? sail_last_error(ctx) : "sail_create failed"); sail_destroy(ctx); return nullptr; }
    napi_value out, h, w, hgt;
    napi_create_object(env, &out);
    napi_create_external(env, ctx, nullptr, nullptr, &h);
    napi_create_int32(env, ci.width, &w);
    napi_create_int32(env, ci.height, &hgt);
    napi_set_named_property(env, out, "handle", h);
    napi_set_named_property(env, out, "width", w);
    napi_set_named_property(env, out, "height", hgt);
    return out;
}
static napi_value Destroy(napi_env env, napi_callback_info info) {
    size_t argc = 1; napi_value a[1];
    NAPI_OK(napi_get_cb_info(env, info, &argc, a, nullptr, nullptr));
    sail_destroy(handle(env, a[0]));
    return nullptr;
}
// configure(handle, cfg)        [setProgram(new TraceShader(...)) / setProgram(new RenderShader(...))]
static napi_value Configure(napi_env env, napi_callback_info info) {
    size_t argc = 2; napi_value a[2];
    NAPI_OK(napi_get_cb_info(env, info, &argc, a, nullptr, nullptr));
    sail_ctx* ctx = handle(env, a[0]);
    sail_config c; sail_default_config(&c);
    c.shape_mask = (uint32_t)numberProp(env, a[1], "shapeMask");
    c.material_mask = (uint32_t)numberProp(env, a[1], "materialMask");
    c.texture_mask = (uint32_t)numberProp(env, a[1], "textureMask");
    c.light_mask = (uint32_t)numberProp(env, a[1], "lightMask");
    c.max_bounces = (int32_t)numberProp(env, a[1], "maxBounces");
    c.accum_mode = (int32_t)numberProp(env, a[1], "accumMode");
    c.kernel = (int32_t)numberProp(env, a[1], "kernel");
    c.filter_id = (int32_t)numberProp(env, a[1], "filterId");
    c.filter_gamma_c = (float)numberProp(env, a[1], "filterGammaC");
    floatArrayProp(env, a[1], "filterRadius", c.filter_radius, 2);
    floatArrayProp(env, a[1], "filterWeights", c.filter_weights, 16);
    check(env, ctx, sail_configure(ctx, &c));
    return nullptr;
}
// uploadScene(handle, objects, texparams, lights)        [Tracer.update]
static napi_value UploadScene(napi_env env, napi_callback_info info) {
    size_t argc = 4; napi_value a[4];
    NAPI_OK(napi_get_cb_info(env, info, &argc, a, nullptr, nullptr));
    sail_ctx* ctx = handle(env, a[0]);
    size_t no, nt, nl;
    float *o = f32(env, a[1], &no), *t = f32(env, a[2], &nt), *l = f32(env, a[3], &nl);
    check(env, ctx, sail_upload_scene(ctx, o, (int32_t)(no / SAIL_OBJECTS_LENGTH), t, (int32_t)(nt / SAIL_TEXPARAMS_LENGTH), l,
                                      (int32_t)(nl / SAIL_LIGHTS_LENGTH)));
    return nullptr;
}
// uploadObjects(handle, objects)        [Tracer.updateObjects]
static napi_value UploadObjects(napi_env env, napi_callback_info info) {
    size_t argc = 2; napi_value a[2];
    NAPI_OK(napi_get_cb_info(env, info, &argc, a, nullptr, nullptr));
    sail_ctx* ctx = handle(env, a[0]);
    size_t no;
    float* o = f32(env, a[1], &no);
    check(env, ctx, sail_upload_objects(ctx, o, (int32_t)(no / SAIL_OBJECTS_LENGTH)));
    return nullptr;
}
// render(handle, matrix16, eye3, textureWeight, timeSinceStart)        [Tracer.render]
static napi_value Render(napi_env env, napi_callback_info info) {
    size_t argc = 5; napi_value a[5];
    NAPI_OK(napi_get_cb_info(env, info, &argc, a, nullptr, nullptr));
    sail_ctx* ctx = handle(env, a[0]);
    size_t nm, ne; double w = 0, time = 0;
    float *m = f32(env, a[1], &nm), *e = f32(env, a[2], &ne);
    napi_get_value_double(env, a[3], &w);
    napi_get_value_double(env, a[4], &time);
    if (nm != 16 || ne != 3) { napi_throw_range_error(env, nullptr, "matrix must have 16 floats, eye 3"); return nullptr; }
    check(env, ctx, sail_render(ctx, m, e, (float)w, (float)time));
    return nullptr;
}
// present(handle) -> Uint8ClampedArray(W*H*4)        [renderShader.render('triangle')]
static napi_value Present(napi_env env, napi_callback_info info) {
    size_t argc = 3; napi_value a[3];
    NAPI_OK(napi_get_cb_info(env, info, &argc, a, nullptr, nullptr));
    sail_ctx* ctx = handle(env, a[0]);
    int32_t W = 0, H = 0;
    napi_get_value_int32(env, a[1], &W);
    napi_get_value_int32(env, a[2], &H);
    void* data = nullptr; napi_value buf, out;
    size_t bytes = (size_t)W * H * 4;
    NAPI_OK(napi_create_arraybuffer(env, bytes, &data, &buf));
    if (!check(env, ctx, sail_present(ctx, nullptr, (uint8_t*)data))) return nullptr;
    NAPI_OK(napi_create_typedarray(env, napi_uint8_clamped_array, bytes, buf, 0, &out));
    return out;
}
// presentBegin(handle, Uint8ClampedArray out) / presentEnd(handle): the display pass as a swapchain of two -- the caller keeps
// `out` alive and does not read it before presentEnd() returns; the next render() may be issued in between.
static napi_value PresentBegin(napi_env env, napi_callback_info info) {
    size_t argc = 2; napi_value a[2];
    NAPI_OK(napi_get_cb_info(env, info, &argc, a, nullptr, nullptr));
    sail_ctx* ctx = handle(env, a[0]);
    napi_typedarray_type type; napi_value buf; size_t off, len; void* data = nullptr;
    NAPI_OK(napi_get_typedarray_info(env, a[1], &type, &len, &data, &buf, &off));
    check(env, ctx, sail_present_begin(ctx, (uint8_t*)data));
    return nullptr;
}
static napi_value PresentEnd(napi_env env, napi_callback_info info) {
    size_t argc = 1; napi_value a[1];
    NAPI_OK(napi_get_cb_info(env, info, &argc, a, nullptr, nullptr));
    check(env, handle(env, a[0]), sail_present_end(handle(env, a[0])));
    return nullptr;
}
// readAov(handle, which, Float32Array|Int32Array out)        [gl.readPixels on frameCache[i]]
static napi_value ReadAov(napi_env env, napi_callback_info info) {
    size_t argc = 3; napi_value a[3];
    NAPI_OK(napi_get_cb_info(env, info, &argc, a, nullptr, nullptr));
    sail_ctx* ctx = handle(env, a[0]);
    int32_t which = 0;
    napi_get_value_int32(env, a[1], &which);
    napi_typedarray_type type; napi_value buf; size_t off, len; void* data = nullptr;
    NAPI_OK(napi_get_typedarray_info(env, a[2], &type, &len, &data, &buf, &off));
    check(env, ctx, sail_read_aov(ctx, which, data, len * 4));
    return nullptr;
}

// renderN(handle, Float32Array frames /* count x 21: matrix16, eye3, textureWeight, timeSinceStart */)   [k x Tracer.render]
static napi_value RenderN(napi_env env, napi_callback_info info) {
    size_t argc = 2; napi_value a[2];
    NAPI_OK(napi_get_cb_info(env, info, &argc, a, nullptr, nullptr));
    sail_ctx* ctx = handle(env, a[0]);
    size_t n = 0;
    float* f = f32(env, a[1], &n);
    static_assert(sizeof(sail_frame) == 21 * sizeof(float), "sail_frame is 21 consecutive floats");
    if (!f || n == 0 || n % 21) { napi_throw_range_error(env, nullptr, "frames must hold 21 floats per frame"); return nullptr; }
    check(env, ctx, sail_render_n(ctx, (const sail_frame*)f, (int32_t)(n / 21)));
    return nullptr;
}
// reset(handle)        [a fresh frameCache]
static napi_value Reset(napi_env env, napi_callback_info info) {
    size_t argc = 1; napi_value a[1];
    NAPI_OK(napi_get_cb_info(env, info, &argc, a, nullptr, nullptr));
    check(env, handle(env, a[0]), sail_reset(handle(env, a[0])));
    return nullptr;
}
// specialiseWait(handle) -> true when the scene-specialised kernel is in place (false: ahead-of-time variant keeps running)
static napi_value SpecialiseWait(napi_env env, napi_callback_info info) {
    size_t argc = 1; napi_value a[1];
    NAPI_OK(napi_get_cb_info(env, info, &argc, a, nullptr, nullptr));
    sail_ctx* ctx = handle(env, a[0]);
    int rc = sail_specialise_wait(ctx);
    if (rc < 0) { check(env, ctx, rc); return nullptr; }
    napi_value out;
    napi_get_boolean(env, rc == 0, &out);
    return out;
}
// kernelInfo(handle) -> string
static napi_value KernelInfo(napi_env env, napi_callback_info info) {
    size_t argc = 1; napi_value a[1];
    NAPI_OK(napi_get_cb_info(env, info, &argc, a, nullptr, nullptr));
    char buf[400];
    int n = sail_kernel_info(handle(env, a[0]), buf, sizeof(buf));
    napi_value out;
    napi_create_string_utf8(env, n < 0 ? "" : buf, NAPI_AUTO_LENGTH, &out);
    return out;
}
// intersectRays(handle, Float32Array rays /* count x 6 */, Int32Array hitId, Float32Array t)        [Pickup.pick, pickup.js:46-63]
static napi_value IntersectRays(napi_env env, napi_callback_info info) {
    size_t argc = 4; napi_value a[4];
    NAPI_OK(napi_get_cb_info(env, info, &argc, a, nullptr, nullptr));
    sail_ctx* ctx = handle(env, a[0]);
    size_t nr = 0, nt = 0;
    float* rays = f32(env, a[1], &nr);
    napi_typedarray_type type; napi_value buf; size_t off, ni = 0; void* ids = nullptr;
    NAPI_OK(napi_get_typedarray_info(env, a[2], &type, &ni, &ids, &buf, &off));
    float* t = f32(env, a[3], &nt);
    if (!rays || nr % 6 || type != napi_int32_array || ni != nr / 6 || nt != nr / 6) { napi_throw_range_error(env, nullptr, "rays: 6 floats each; hitId (Int32Array), t: one element per ray"); return nullptr; }
    check(env, ctx, sail_intersect_rays(ctx, rays, (int32_t)(nr / 6), (int32_t*)ids, t, nullptr, nullptr));
    return nullptr;
}
// groupId() -> Uint8Array(128); groupJoin(handle, nranks, rank, id, shard, bandRows): one GPU per process (a cluster of Node
// workers), images arrive on rank 0 (include/sail_b200.h sail_group_join)
static napi_value GroupId(napi_env env, napi_callback_info info) {
    void* data = nullptr; napi_value buf, out;
    NAPI_OK(napi_create_arraybuffer(env, SAIL_GROUP_ID_BYTES, &data, &buf));
    if (sail_group_id((uint8_t*)data) != SAIL_OK) { napi_throw_error(env, nullptr, "sail_group_id failed (NCCL not available?)"); return nullptr; }
    NAPI_OK(napi_create_typedarray(env, napi_uint8_array, SAIL_GROUP_ID_BYTES, buf, 0, &out));
    return out;
}
static napi_value GroupJoin(napi_env env, napi_callback_info info) {
    size_t argc = 6; napi_value a[6];
    NAPI_OK(napi_get_cb_info(env, info, &argc, a, nullptr, nullptr));
    sail_ctx* ctx = handle(env, a[0]);
    int32_t nranks = 0, rank = 0, shard = 0, band = 0;
    napi_get_value_int32(env, a[1], &nranks);
    napi_get_value_int32(env, a[2], &rank);
    napi_typedarray_type type; napi_value buf; size_t off, len = 0; void* id = nullptr;
    NAPI_OK(napi_get_typedarray_info(env, a[3], &type, &len, &id, &buf, &off));
    napi_get_value_int32(env, a[4], &shard);
    if (argc > 5) napi_get_value_int32(env, a[5], &band);
    if (type != napi_uint8_array || len != SAIL_GROUP_ID_BYTES) { napi_throw_range_error(env, nullptr, "group id: the Uint8Array(128) of groupId()"); return nullptr; }
    check(env, ctx, sail_group_join(ctx, nranks, rank, (const uint8_t*)id, shard, band, SAIL_COMBINE_AUTO));
    return nullptr;
}

static napi_value Init(napi_env env, napi_value exports) {
    napi_property_descriptor d[] = {
        {"create", nullptr, Create, nullptr, nullptr, nullptr, napi_default, nullptr},
        {"destroy", nullptr, Destroy, nullptr, nullptr, nullptr, napi_default, nullptr},
        {"configure", nullptr, Configure, nullptr, nullptr, nullptr, napi_default, nullptr},
        {"uploadScene", nullptr, UploadScene, nullptr, nullptr, nullptr, napi_default, nullptr},
        {"uploadObjects", nullptr, UploadObjects, nullptr, nullptr, nullptr, napi_default, nullptr},
        {"render", nullptr, Render, nullptr, nullptr, nullptr, napi_default, nullptr},
        {"renderN", nullptr, RenderN, nullptr, nullptr, nullptr, napi_default, nullptr},
        {"reset", nullptr, Reset, nullptr, nullptr, nullptr, napi_default, nullptr},
        {"present", nullptr, Present, nullptr, nullptr, nullptr, napi_default, nullptr},
        {"presentBegin", nullptr, PresentBegin, nullptr, nullptr, nullptr, napi_default, nullptr},
        {"presentEnd", nullptr, PresentEnd, nullptr, nullptr, nullptr, napi_default, nullptr},
        {"readAov", nullptr, ReadAov, nullptr, nullptr, nullptr, napi_default, nullptr},
        {"specialiseWait", nullptr, SpecialiseWait, nullptr, nullptr, nullptr, napi_default, nullptr},
        {"kernelInfo", nullptr, KernelInfo, nullptr, nullptr, nullptr, napi_default, nullptr},
        {"intersectRays", nullptr, IntersectRays, nullptr, nullptr, nullptr, napi_default, nullptr},
        {"groupId", nullptr, GroupId, nullptr, nullptr, nullptr, napi_default, nullptr},
        {"groupJoin", nullptr, GroupJoin, nullptr, nullptr, nullptr, napi_default, nullptr},
    };
    napi_define_properties(env, exports, sizeof(d) / sizeof(d[0]), d);
    return exports;
}
NAPI_MODULE(NODE_GYP_MODULE_NAME, Init)
