// sail_mega_variants.cu -- instantiates the megakernel for the four feature sets F = 4*SAIL_VGROUP .. +3
// (compiled once per group so the 16 variants build in parallel).  See sail_shapes.cuh "Kernel specialisation".
#include <cuda_runtime.h>
#include "sail_mega.cuh"
#include "sail_wavefront.cuh"

#ifndef SAIL_VGROUP
#error "compile with -DSAIL_VGROUP=0..3"
#endif
#define SAIL_CAT2(a, b) a##b
#define SAIL_CAT(a, b) SAIL_CAT2(a, b)

extern "C" cudaError_t SAIL_CAT(sail_launch_mega_g, SAIL_VGROUP)(int F, const SailLaunch* L, dim3 grid, cudaStream_t stream) {
    switch (F - 4 * SAIL_VGROUP) {
        case 0: sail_trace_megakernel<4 * SAIL_VGROUP + 0><<<grid, SAIL_CTA_THREADS, L->sceneSmem, stream>>>(*L); break;
        case 1: sail_trace_megakernel<4 * SAIL_VGROUP + 1><<<grid, SAIL_CTA_THREADS, L->sceneSmem, stream>>>(*L); break;
        case 2: sail_trace_megakernel<4 * SAIL_VGROUP + 2><<<grid, SAIL_CTA_THREADS, L->sceneSmem, stream>>>(*L); break;
        case 3: sail_trace_megakernel<4 * SAIL_VGROUP + 3><<<grid, SAIL_CTA_THREADS, L->sceneSmem, stream>>>(*L); break;
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

extern "C" cudaError_t SAIL_CAT(sail_launch_wave_g, SAIL_VGROUP)(int F, const SailLaunch* L, dim3 grid, cudaStream_t stream) {
    switch (F - 4 * SAIL_VGROUP) {
        case 0: sail_trace_wavefront<4 * SAIL_VGROUP + 0><<<grid, SAIL_CTA_THREADS, L->sceneSmem, stream>>>(*L); break;
        case 1: sail_trace_wavefront<4 * SAIL_VGROUP + 1><<<grid, SAIL_CTA_THREADS, L->sceneSmem, stream>>>(*L); break;
        case 2: sail_trace_wavefront<4 * SAIL_VGROUP + 2><<<grid, SAIL_CTA_THREADS, L->sceneSmem, stream>>>(*L); break;
        case 3: sail_trace_wavefront<4 * SAIL_VGROUP + 3><<<grid, SAIL_CTA_THREADS, L->sceneSmem, stream>>>(*L); break;
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}
