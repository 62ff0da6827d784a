// sail_jit.h -- scene-specialised kernels compiled at run time (see sail_jit.cpp).
#ifndef SAIL_JIT_H
#define SAIL_JIT_H
#include "sail_scene.h"

#define SAIL_JIT_MAX_RUNS 16  // runs of equal shape category; beyond this the ahead-of-time variants take over

struct SailJitKernel {
    void* fn;        // CUfunction (null after a compile-only request)
    char info[200];  // "jit mega<F=0> n=3 shapes=1,9,2 (nvrtc 12.9, 0.80 s)" or "... (cached)"
    unsigned block;  // threads per CTA the kernel was compiled for
    void* stage[5];  // kind 2 (sail_queue.cuh): raygen, scan, scatter, shade, accumulate; fn = stage[0]
};
#define SAIL_JIT_WF_THREADS_DEFAULT 128  // paths per wavefront CTA of the run-time compiled kernel (SAIL_B200_WF_THREADS overrides)
#define SAIL_JIT_CTA_THREADS 128u  // = SAIL_CTA_THREADS of sail_mega.cuh

// A compilation in flight or finished; lives for the process (cubins are cached per scene structure and shared by all devices).
struct SailJitJob;

// kind: 0 megakernel, 1 shared-memory wavefront, 2 queue wavefront (five stage kernels in one module).  Starts (or finds) the NVRTC compilation of the kernel specialised for `scene`'s structure on a
// worker thread and returns its job, or null with a message in `err` (NVRTC missing, scene not specialisable ...): the caller then
// keeps using the ahead-of-time variant.  Never blocks on the compiler.
// rngSmall bit 1: Russian roulette compiled in (sail_config.rr_start_bounce > 0).  rngSmall bit 0: the caller guarantees |argument| < 2^22 for every sine of the hash RNG (sail_shade.cuh sinPair), so the
// large-argument reduction is not compiled.
SailJitJob* sail_jit_request(int kind, int F, int minCtas, int rngSmall, const SailSceneBlock* scene, char* err, int errlen);
// 0: compiled, module loaded on `device` (the current device), `out` filled.  1: still compiling (only when wait == 0).
// < 0: failed, message in `err`.  device < 0: compile only (needs neither the driver library nor a GPU), out->fn stays null.
int sail_jit_poll(SailJitJob* job, int device, int wait, SailJitKernel* out, char* err, int errlen);
// Launches `fn` with the SailLaunch parameter block at `launch` on `stream`.  Returns 0 or a CUDA driver error code.
int sail_jit_launch(void* fn, const void* launch, unsigned gridX, unsigned gridY, unsigned block, unsigned smemBytes, void* stream,
                    char* err, int errlen);
#endif
