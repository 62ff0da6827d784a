// sail_wavefront.cuh -- wavefront (per-bounce kernels over compacted ray queues) variant.  Placeholder
// until the queue kernels land: selecting SAIL_KERNEL_WAVEFRONT fails loudly.
#ifndef SAIL_WAVEFRONT_CUH
#define SAIL_WAVEFRONT_CUH
#include "sail_mega.cuh"
struct SailWavefront {
    cudaError_t render(const SailLaunch&, cudaStream_t, int* launched) { *launched = 0; return cudaErrorNotSupported; }
    void release() {}
};
#endif
