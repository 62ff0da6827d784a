// sail_jit.cpp -- per-scene kernel specialisation at run time.
//
// The reference assembles its fragment shader per scene: only the plugins the scene uses, with the scene's object
// count baked in (/root/reference/src/shader/generator.js:107-123, src/scene/scene.js:70-112, src/core/tracer.js:43).
// The 16 ahead-of-time variants (sail_mega_variants.cu) reproduce the "only what is used" half.  This file adds the
// other half: NVRTC compiles the same kernel body with the object count and the shape category of every object slot
// as compile-time constants, so the nearest-hit scan becomes straight-line code (sail_shapes.cuh, SAIL_JIT_NRUNS).
// Scene *values* (positions, sizes, materials) stay in the shared-memory scene block, so dragging an object
// (Renderer.updateObjects) never recompiles; only a change of structure does, and kernels are cached per process.
//
// NVRTC and the CUDA driver library are opened with dlopen: the library links against neither, and when one of them is
// missing sail_jit_get() reports it and the caller runs the ahead-of-time variant (still a CUDA kernel -- there is no
// CPU path anywhere).
#include "sail_jit.h"

#include <dlfcn.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <chrono>
#include <condition_variable>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "sail_jit_sources.inc"

namespace {

typedef int (*nvrtcCreateProgram_t)(void**, const char*, const char*, int, const char* const*, const char* const*);
typedef int (*nvrtcCompileProgram_t)(void*, int, const char* const*);
typedef int (*nvrtcGetSize_t)(void*, size_t*);
typedef int (*nvrtcGetData_t)(void*, char*);
typedef int (*nvrtcDestroyProgram_t)(void**);
typedef int (*nvrtcVersion_t)(int*, int*);
typedef int (*cuModuleLoadData_t)(void**, const void*);
typedef int (*cuModuleGetFunction_t)(void**, void*, const char*);
typedef int (*cuLaunchKernel_t)(void*, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, void*, void**, void**);
typedef int (*cuGetErrorString_t)(int, const char**);
typedef int (*cuFuncSetAttribute_t)(void*, int, int);

struct Api {
    bool tried = false, ok = false;        // NVRTC
    bool cuTried = false, cuOk = false;    // driver library (needed to load and launch, not to compile)
    std::string why, cuWhy;
    int nvrtcMajor = 0, nvrtcMinor = 0;
    nvrtcCreateProgram_t createProgram = nullptr;
    nvrtcCompileProgram_t compileProgram = nullptr;
    nvrtcGetSize_t logSize = nullptr, cubinSize = nullptr;
    nvrtcGetData_t log = nullptr, cubin = nullptr;
    nvrtcDestroyProgram_t destroyProgram = nullptr;
    cuModuleLoadData_t moduleLoadData = nullptr;
    cuModuleGetFunction_t moduleGetFunction = nullptr;
    cuLaunchKernel_t launchKernel = nullptr;
    cuGetErrorString_t errorString = nullptr;
    cuFuncSetAttribute_t funcSetAttribute = nullptr;
};
Api g_api;
std::mutex g_mutex;

void* openFirst(const std::vector<std::string>& names, std::string* tried) {
    for (const std::string& n : names) {
        if (n.empty()) continue;
        if (void* h = dlopen(n.c_str(), RTLD_NOW | RTLD_LOCAL)) return h;
        *tried += n + " ";
    }
    return nullptr;
}

void waitForCompilersAtExit();

bool loadApi() {
    if (g_api.tried) return g_api.ok;
    g_api.tried = true;
    const char* off = getenv("SAIL_B200_JIT");
    if (off && off[0] == '0') { g_api.why = "disabled by SAIL_B200_JIT=0"; return false; }
    std::string tried;
    const char* envPath = getenv("SAIL_B200_NVRTC");
    void* nv = openFirst({envPath ? envPath : "", "/usr/local/cuda/lib64/libnvrtc.so.12", "libnvrtc.so.12", "libnvrtc.so"}, &tried);
    if (!nv) { g_api.why = "NVRTC not found (tried " + tried + ")"; return false; }
#define SYM(lib, field, name, type) g_api.field = (type)dlsym(lib, name); if (!g_api.field) { g_api.why = std::string("missing symbol ") + name; return false; }
    SYM(nv, createProgram, "nvrtcCreateProgram", nvrtcCreateProgram_t)
    SYM(nv, compileProgram, "nvrtcCompileProgram", nvrtcCompileProgram_t)
    SYM(nv, logSize, "nvrtcGetProgramLogSize", nvrtcGetSize_t)
    SYM(nv, log, "nvrtcGetProgramLog", nvrtcGetData_t)
    SYM(nv, cubinSize, "nvrtcGetCUBINSize", nvrtcGetSize_t)
    SYM(nv, cubin, "nvrtcGetCUBIN", nvrtcGetData_t)
    SYM(nv, destroyProgram, "nvrtcDestroyProgram", nvrtcDestroyProgram_t)
#undef SYM
    if (nvrtcVersion_t ver = (nvrtcVersion_t)dlsym(nv, "nvrtcVersion")) ver(&g_api.nvrtcMajor, &g_api.nvrtcMinor);
    // Registered AFTER libnvrtc was loaded, so at exit it runs BEFORE libnvrtc's own finalisers: a compilation still in flight is
    // given a bounded time to finish instead of racing the teardown of the compiler it is running in.
    atexit(waitForCompilersAtExit);
    g_api.ok = true;
    return true;
}

bool loadDriver() {
    if (g_api.cuTried) return g_api.cuOk;
    g_api.cuTried = true;
    std::string tried;
    void* cu = openFirst({"libcuda.so.1", "libcuda.so"}, &tried);
    if (!cu) { g_api.cuWhy = "CUDA driver library not found"; return false; }
#define SYM(lib, field, name, type) g_api.field = (type)dlsym(lib, name); if (!g_api.field) { g_api.cuWhy = std::string("missing symbol ") + name; return false; }
    SYM(cu, moduleLoadData, "cuModuleLoadData", cuModuleLoadData_t)
    SYM(cu, moduleGetFunction, "cuModuleGetFunction", cuModuleGetFunction_t)
    SYM(cu, launchKernel, "cuLaunchKernel", cuLaunchKernel_t)
    SYM(cu, errorString, "cuGetErrorString", cuGetErrorString_t)
    SYM(cu, funcSetAttribute, "cuFuncSetAttribute", cuFuncSetAttribute_t)
#undef SYM
    g_api.cuOk = true;
    return true;
}

void setErr(char* err, int errlen, const std::string& s) {
    if (err && errlen > 0) snprintf(err, (size_t)errlen, "%s", s.c_str());
}


struct Plan {
    std::string key, src, head, cats, masks;
    unsigned block = SAIL_JIT_CTA_THREADS;
    int kind = 0, F = 0, n = 0;
};

}  // namespace

struct SailJitJob {
    Plan plan;
    std::mutex m;
    std::condition_variable cv;
    bool done = false;
    int rc = 0;  // 0 ok, else failure (message in err)
    std::string err;
    std::vector<char> cubin;
    double secs = 0.0;
    std::map<int, void*> fn;  // device -> CUfunction (guarded by g_mutex)
    std::map<int, void* [5]> stage;  // kind 2: device -> the five stage kernels
};

namespace {

// Jobs live for the process: a scene structure is compiled once and its cubin is loaded on every device that asks.  The
// registry and the jobs are never destroyed and the compiler threads are detached: a host that exits while NVRTC is still
// working must not wait for it -- joining from a static destructor deadlocks, because libnvrtc (opened later, so finalised
// earlier) has already torn down the state its compiler thread is using.  Process exit ends the thread.
struct Registry {
    std::map<std::string, SailJitJob*> jobs;
};
Registry& registry() {
    static Registry* r = new Registry;  // intentionally leaked
    return *r;
}

// Fills the plan (cache key, generated source, launch shape) for `scene`; 0 or an error code with a message.
int makePlan(int kind, int F, int minCtas, int rngSmall, const SailSceneBlock* scene, Plan* plan, char* err, int errlen) {
    const int n = scene->n;
    if (n < 1) { setErr(err, errlen, "empty scene"); return 2; }
    if (kind < 0 || kind > 2) { setErr(err, errlen, "unknown kernel kind"); return 2; }
    // runs of equal shape category in slot order
    std::string cats, runCat, runStart, runLen;
    int nruns = 0;
    for (int i = 0; i < n;) {
        int j = i;
        while (j < n && scene->shape[j].category == scene->shape[i].category) j++;
        const char* sep = nruns ? "," : "";
        runCat += sep + std::to_string(scene->shape[i].category);
        runStart += sep + std::to_string(i);
        runLen += sep + std::to_string(j - i);
        cats += sep + std::to_string(scene->shape[i].category) + (j - i > 1 ? "x" + std::to_string(j - i) : "");
        nruns++;
        i = j;
    }
    if (nruns > SAIL_JIT_MAX_RUNS) { setErr(err, errlen, "more than 16 runs of distinct shapes"); return 2; }
    // which plugins the scene contains (category ids as bit positions; ids outside 0..31 select no arm anyway)
    unsigned shapes = 0, mats = 0, texs = 0, lights = 0, glass = 0, matte = 0, tr = 0;
    auto bit = [](int c) { return (c >= 0 && c < 32) ? (1u << c) : 0u; };
    for (int i = 0; i < n; i++) {
        shapes |= bit(scene->shape[i].category);
        mats |= bit(scene->shape[i].matCategory) | bit(scene->surf[i].matCategory);
        texs |= bit(scene->surf[i].texCategory);
        if (scene->shape[i].matCategory == SAIL_MATTE || scene->surf[i].matCategory == SAIL_MATTE)
            matte |= (scene->surf[i].m[SM_MATTE_SIGMA] < 1e-5f) ? 1u : 2u;  // the shader's test, matte.glsl:13 / :30 (EPSILON)
        if (scene->shape[i].matCategory == SAIL_GLASS || scene->surf[i].matCategory == SAIL_GLASS) {
            const float ur = scene->surf[i].m[SM_GLASS_UR], vr = scene->surf[i].m[SM_GLASS_VR];
            glass |= (ur < 1e-5f && vr < 1e-5f) ? 1u : 2u;  // the shader's test, glass.glsl:13 (EPSILON, define.glsl)
            if (!(ur < 1e-5f && vr < 1e-5f)) tr |= (ur == vr) ? 1u : 2u;  // microfacet.glsl:44 `alphax == alphay`
        }
        if (scene->shape[i].matCategory == SAIL_METAL || scene->surf[i].matCategory == SAIL_METAL)
            tr |= (scene->surf[i].m[SM_METAL_UR] == scene->surf[i].m[SM_METAL_VR]) ? 1u : 2u;
    }
    for (int k = 0; scene->ln > 0 && k <= scene->ln; k++) lights |= bit(scene->light[k].category);
    // an area light samples its geometry: the sampler arms of those shapes are part of the structure too
    unsigned lightShapes = 0;
    for (int k = 0; scene->ln > 0 && k <= scene->ln; k++)
        if (scene->light[k].category == SAIL_AREA && scene->light[k].object >= 0 && scene->light[k].object < n)
            lightShapes |= bit(scene->shape[scene->light[k].object].category);
    // SAIL_B200_JIT_DEFINES="NAME=VALUE,NAME2=VALUE2": extra #defines for A/B runs of experimental code paths
    std::string extraDefs;
    if (const char* e = getenv("SAIL_B200_JIT_DEFINES")) {
        std::string cur;
        for (const char* p = e;; p++) {
            if (*p == ',' || *p == 0) {
                if (!cur.empty()) {
                    const size_t eq = cur.find('=');
                    extraDefs += "#define " + (eq == std::string::npos ? cur + " 1" : cur.substr(0, eq) + " " + cur.substr(eq + 1)) + "\n";
                }
                cur.clear();
                if (*p == 0) break;
            } else cur += *p;
        }
    }
    char masks[200];
    snprintf(masks, sizeof(masks), " plugins=%x/%x/%x/%x/%x/%x/%x/%x", shapes, mats, texs, lights, glass, matte, tr, lightShapes);
    if (!(rngSmall & 1)) snprintf(masks + strlen(masks), sizeof(masks) - strlen(masks), " rng=full-range");
    if (rngSmall & 2) snprintf(masks + strlen(masks), sizeof(masks) - strlen(masks), " +rr");
    if (const char* e = extraDefs.empty() ? nullptr : getenv("SAIL_B200_JIT_DEFINES")) snprintf(masks + strlen(masks), sizeof(masks) - strlen(masks), " -D%.60s", e);
    unsigned block = SAIL_JIT_CTA_THREADS;
    if (kind == 1) {  // wavefront: paths per CTA (sail_wavefront.cuh SAIL_WF_THREADS), 64 registers per thread fill the SM
        // measured (tools/wf_sweep.sh, Mpaths/s at 128 / 256 / 512 paths): config 3 (quadrics) 1744 / 2059 / 2269, config 5 1204 / 1303 / 1336,
        // config 2 3930 / 3956 / 3768
        int wf = (F & 1 /* SAIL_F_QUADRICS */) ? 512 : ((F & 2 /* SAIL_F_GLOSSY */) ? 256 : SAIL_JIT_WF_THREADS_DEFAULT);
        if (const char* e = getenv("SAIL_B200_WF_THREADS")) wf = atoi(e);
        if (wf != 128 && wf != 256 && wf != 384 && wf != 512) { setErr(err, errlen, "SAIL_B200_WF_THREADS must be 128, 256, 384 or 512"); return 2; }
        block = (unsigned)wf;
        if (wf != 128) minCtas = (int)(65536u / (64u * block));
    }
    if (const char* e = getenv("SAIL_B200_MIN_CTAS")) {  // occupancy sweeps (tools/ab_defs.sh)
        const int v = atoi(e);
        if (v >= 1 && v <= 16) minCtas = v;
    }
    char head[96];
    if (kind == 1 && block != 128) snprintf(head, sizeof(head), "jit wavefront<F=%d,%u paths> n=%d shapes=", F, block, n);
    else snprintf(head, sizeof(head), "jit %s<F=%d> n=%d shapes=", kind == 2 ? "queue" : (kind ? "wavefront" : "mega"), F, n);
    // One shared sqrt instead of ~50 inlined copies: pays for quadric kernels that overflow the 32 KB instruction cache (megakernel,
    // 128-path wavefront: config 3 +5..16 %), costs 3.7 % once the 512-path wavefront has taken that pressure away (2318 vs 2404
    // Mpaths/s); metal/glass-only kernels are faster inlined.  SAIL_B200_INLINE_SQRT=0/1 overrides for A/B runs.
    bool outlineSqrt = (F & 1 /* SAIL_F_QUADRICS */) && !(kind == 1 && block >= 256);
    if (const char* e = getenv("SAIL_B200_INLINE_SQRT")) outlineSqrt = e[0] == '0';
    std::string tuning = outlineSqrt ? "#define SAIL_OUTLINE_SQRT 1\n" : "#define SAIL_INLINE_TRANSCENDENTALS 1\n";
    // small scenes without quadrics: the nearest-hit scan inlined at both call sites (path ray, shadow ray) instead of one shared
    // out-of-line copy -- config 5 +2.8 %, config 2 +-0; with quadrics the scan is too large (config 3 -23 %)
    if (!(F & 1 /* SAIL_F_QUADRICS */) && n <= 8) tuning += "#define SAIL_INLINE_SCAN 1\n";
    if (rngSmall & 1) tuning += "#define SAIL_RNG_SMALL_ARGS 1\n";
    tuning += (rngSmall & 2) ? "#define SAIL_JIT_RR 1\n" : "#define SAIL_JIT_RR 0\n";
    char src[6144];
    int len = snprintf(src, sizeof(src),
             "%s#define SAIL_JIT_NRUNS %d\n#define SAIL_JIT_RUN_CATS %s\n#define SAIL_JIT_RUN_STARTS %s\n#define SAIL_JIT_RUN_LENS %s\n"
             "#define SAIL_JIT_SHAPES 0x%xu\n#define SAIL_JIT_MATS 0x%xu\n#define SAIL_JIT_TEXS 0x%xu\n#define SAIL_JIT_LIGHTS 0x%xu\n"
             "#define SAIL_JIT_GLASS %u\n#define SAIL_JIT_MATTE %u\n#define SAIL_JIT_TR %u\n#define SAIL_JIT_LIGHT_SHAPES 0x%xu\n#define SAIL_WF_THREADS %u\n%s"
             "#include \"sail_mega.cuh\"\n#include \"sail_wavefront.cuh\"\n#include \"sail_queue.cuh\"\n",
             extraDefs.c_str(), nruns, runCat.c_str(), runStart.c_str(), runLen.c_str(), shapes, mats, texs, lights, glass, matte, tr, lightShapes,
             kind == 1 ? block : 128u,
             // kernels that are not instruction-cache bound also inline the sines of the RNG, the sampler's sincos and the sphere's
             // acos / sin (measured: config 2 +2.2 %, config 5 +2.9 %, config 3 +0.5 %)
             tuning.c_str());
    if (kind == 2) {  // one entry point per stage of sail_queue.cuh; each stage gets its own register budget
        static const char* const stages[5] = {"Raygen", "Scan", "Scatter", "Shade", "Accumulate"};
        for (int st = 0; st < 5; st++)
            len += snprintf(src + len, sizeof(src) - (size_t)len,
                            "extern \"C\" __global__ void __launch_bounds__(%u, %d) sail_jit_q%d(const __grid_constant__ SailQLaunch L) { sailQ%s<%d>(L); }\n",
                            block, st == 3 ? minCtas : (st == 1 ? minCtas : 8), st, stages[st], F);
    } else {
        snprintf(src + len, sizeof(src) - (size_t)len,
                 "extern \"C\" __global__ void __launch_bounds__(%u, %d) sail_jit_kernel(const __grid_constant__ SailLaunch L) { %s<%d>(L); }\n",
                 block, minCtas, kind ? "sailWaveBody" : "sailMegaBody", F);
    }
    plan->src = src;
    // The key is the generated source itself: whatever is baked into the kernel decides which cached kernel may be reused.
    plan->key = plan->src;
    plan->head = head;
    plan->cats = cats;
    plan->masks = masks;
    plan->block = block;
    plan->kind = kind;
    plan->F = F;
    plan->n = n;
    return 0;
}

void compileJob(SailJitJob* job) {
    void* prog = nullptr;
    int rc = 0;
    std::string msg;
    std::vector<char> cubin;
    const auto t0 = std::chrono::steady_clock::now();
    if (g_api.createProgram(&prog, job->plan.src.c_str(), "sail_jit.cu", kSailJitCount, kSailJitTexts, kSailJitNames) != 0) {
        rc = 3;
        msg = "nvrtcCreateProgram failed";
    } else {
        const char* opts[] = {"--gpu-architecture=sm_100a", "--fmad=false", "-std=c++17", "-lineinfo"};
        if (g_api.compileProgram(prog, 4, opts) != 0) {
            size_t ls = 0;
            g_api.logSize(prog, &ls);
            std::string log(ls ? ls : 1, '\0');
            if (ls) g_api.log(prog, &log[0]);
            rc = 4;
            msg = "nvrtcCompileProgram failed: " + log.substr(0, 400);
        } else {
            size_t cs = 0;
            g_api.cubinSize(prog, &cs);
            cubin.resize(cs);
            g_api.cubin(prog, cubin.data());
        }
        g_api.destroyProgram(&prog);
    }
    const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (rc == 0)
        if (const char* dump = getenv("SAIL_B200_JIT_DUMP")) {  // keep the cubin for cuobjdump / nvdisasm / tools/sass_lines.py
            std::string name = std::string(dump) + "/sail_jit_" + (job->plan.kind == 2 ? "queue" : (job->plan.kind ? "wave" : "mega")) + "_F" + std::to_string(job->plan.F) + "_n" + std::to_string(job->plan.n) + ".cubin";
            if (FILE* f = fopen(name.c_str(), "wb")) { fwrite(cubin.data(), 1, cubin.size(), f); fclose(f); }
        }
    {
        std::lock_guard<std::mutex> lk(job->m);
        job->rc = rc;
        job->err = msg;
        job->cubin.swap(cubin);
        job->secs = secs;
        job->done = true;
    }
    job->cv.notify_all();
}

void waitForCompilersAtExit() {
    std::vector<SailJitJob*> jobs;
    {
        std::lock_guard<std::mutex> lock(g_mutex);
        for (auto& kv : registry().jobs) jobs.push_back(kv.second);
    }
    const auto deadline = std::chrono::steady_clock::now() + std::chrono::seconds(20);
    for (SailJitJob* job : jobs) {
        std::unique_lock<std::mutex> lk(job->m);
        job->cv.wait_until(lk, deadline, [job] { return job->done; });
    }
}

}  // namespace

SailJitJob* sail_jit_request(int kind, int F, int minCtas, int rngSmall, const SailSceneBlock* scene, char* err, int errlen) {
    std::lock_guard<std::mutex> lock(g_mutex);
    if (!loadApi()) { setErr(err, errlen, g_api.why); return nullptr; }
    Plan plan;
    if (makePlan(kind, F, minCtas, rngSmall, scene, &plan, err, errlen) != 0) return nullptr;
    auto it = registry().jobs.find(plan.key);
    if (it != registry().jobs.end()) return it->second;
    SailJitJob* job = new SailJitJob;
    job->plan = plan;
    registry().jobs[plan.key] = job;
    std::thread(compileJob, job).detach();
    return job;
}

int sail_jit_poll(SailJitJob* job, int device, int wait, SailJitKernel* out, char* err, int errlen) {
    if (!job) { setErr(err, errlen, "no job"); return -1; }
    {
        std::unique_lock<std::mutex> lk(job->m);
        if (!job->done) {
            if (!wait) return 1;
            job->cv.wait(lk, [job] { return job->done; });
        }
        if (job->rc != 0) { setErr(err, errlen, job->err); return -job->rc; }
    }
    const Plan& p = job->plan;
    out->block = p.block;
    out->fn = nullptr;
    if (device < 0) {
        snprintf(out->info, sizeof(out->info), "%s%s%s (nvrtc %d.%d, %.2f s, %zu byte cubin, not loaded)", p.head.c_str(), p.cats.c_str(), p.masks.c_str(),
                 g_api.nvrtcMajor, g_api.nvrtcMinor, job->secs, job->cubin.size());
        return 0;
    }
    std::lock_guard<std::mutex> lock(g_mutex);
    if (!loadDriver()) { setErr(err, errlen, g_api.cuWhy); return -1; }
    auto it = job->fn.find(device);
    if (it == job->fn.end()) {
        void *mod = nullptr, *fn = nullptr;
        int cr = g_api.moduleLoadData(&mod, job->cubin.data());
        if (cr == 0 && p.kind == 2) {
            for (int st = 0; st < 5 && cr == 0; st++) {
                char name[32];
                snprintf(name, sizeof(name), "sail_jit_q%d", st);
                cr = g_api.moduleGetFunction(&job->stage[device][st], mod, name);
            }
            fn = job->stage[device][0];
        } else if (cr == 0) cr = g_api.moduleGetFunction(&fn, mod, "sail_jit_kernel");
        // A 512-path wavefront holds 45 KB of static shared memory (exchange slots); together with the scene block of a large
        // scene (up to 17 KB, dynamic) that passes the 48 KB a kernel gets without opting in.
        if (cr == 0 && p.block > 128) cr = g_api.funcSetAttribute(fn, 8 /* CU_FUNC_ATTRIBUTE_MAX_DYNAMIC_SHARED_SIZE_BYTES */, (int)sizeof(SailSceneBlock));
        if (cr != 0) {
            const char* es = nullptr;
            g_api.errorString(cr, &es);
            setErr(err, errlen, std::string("loading the compiled kernel failed: ") + (es ? es : "?"));
            return -5;
        }
        it = job->fn.emplace(device, fn).first;
    }
    out->fn = it->second;
    for (int st = 0; st < 5; st++) out->stage[st] = p.kind == 2 ? job->stage[device][st] : nullptr;
    snprintf(out->info, sizeof(out->info), "%s%s%s (nvrtc %d.%d, %.2f s)", p.head.c_str(), p.cats.c_str(), p.masks.c_str(), g_api.nvrtcMajor, g_api.nvrtcMinor, job->secs);
    return 0;
}

int sail_jit_launch(void* fn, const void* launch, unsigned gridX, unsigned gridY, unsigned block, unsigned smemBytes, void* stream,
                    char* err, int errlen) {
    void* params[1] = {const_cast<void*>(launch)};
    const int cr = g_api.launchKernel(fn, gridX, gridY, 1, block, 1, 1, smemBytes, stream, params, nullptr);
    if (cr != 0) {
        const char* es = nullptr;
        g_api.errorString(cr, &es);
        setErr(err, errlen, std::string("cuLaunchKernel: ") + (es ? es : "?"));
    }
    return cr;
}
