// sail_jit.h -- scene-specialised kernels compiled at run time (see sail_jit.cpp).
#ifndef SAIL_JIT_H
#define SAIL_JIT_H
#include "sail_scene.h"

#define SAIL_JIT_MAX_RUNS 16  // runs of equal shape category; beyond this the ahead-of-time variants take over

struct SailJitKernel {
    void* fn;        // CUfunction
    char info[200];  // "jit mega<F=0> n=3 shapes=1,9,2 (nvrtc 12.9, 0.80 s)" or "... (cached)"
    unsigned block;  // threads per CTA the kernel was compiled for
    int pathWarps;   // kind 2: path warps per CTA (sail_server.cuh); 0 otherwise
};
#define SAIL_JIT_WF_THREADS_DEFAULT 128  // paths per wavefront CTA of the run-time compiled kernel (SAIL_B200_WF_THREADS overrides)
#define SAIL_JIT_CTA_THREADS 128u  // = SAIL_CTA_THREADS of sail_mega.cuh (kinds 0 and 1)

// kind: 0 megakernel, 1 wavefront, 2 megakernel with a server warp (sail_server.cuh; run-time compiled only).  Returns 0 and fills `out`, or non-zero with a message in `err` (NVRTC or the
// driver library missing, compilation failed, scene too large ...): the caller then uses the ahead-of-time variant.
// device < 0 compiles only (NVRTC needs neither the driver library nor a GPU) and leaves out->fn null.
int sail_jit_get(int device, int kind, int F, int minCtas, const SailSceneBlock* scene, SailJitKernel* out, char* err, int errlen);
// Launches `fn` with the SailLaunch parameter block at `launch` on `stream`.  Returns 0 or a CUDA driver error code.
int sail_jit_launch(void* fn, const void* launch, unsigned gridX, unsigned gridY, unsigned block, unsigned smemBytes, void* stream,
                    char* err, int errlen);
#endif
