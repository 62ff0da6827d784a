// Micro-benchmark of the display kernels on a synthetic 1080p image: nvcc ... -DSAIL_PRESENT_MIN_CTAS\(K\)=N tools/ubench/present_win.cu
// (occupancy sweeps of sail_present_kernel<1|2> without rebuilding the library).  Prints ms per launch, events on the stream.
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
#include "../../sail_b200/csrc/sail_present.cuh"
int main(int argc, char** argv) {
    const int W = 1920, H = 1080, filter = argc > 1 ? atoi(argv[1]) : 2;
    const size_t n = (size_t)W * H;
    std::vector<float> h(n * 3);
    unsigned s = 12345u;
    for (auto& v : h) { s = s * 1664525u + 1013904223u; v = (s >> 8) * (1.0f / 16777216.0f); }
    float *color, *pos, *nrm; float4 *pad, *pad2; unsigned char* out;
    cudaMalloc(&color, n * 12); cudaMalloc(&pos, n * 12); cudaMalloc(&nrm, n * 12); cudaMalloc(&pad, n * 16); cudaMalloc(&pad2, n * 16); cudaMalloc(&out, n * 4);
    cudaMemcpy(color, h.data(), n * 12, cudaMemcpyHostToDevice); cudaMemcpy(pos, h.data(), n * 12, cudaMemcpyHostToDevice); cudaMemcpy(nrm, h.data(), n * 12, cudaMemcpyHostToDevice);
    PresentParams P{};
    P.W = W; P.H = H; P.filter_id = filter; P.gamma_c = 2.2f; P.radius[0] = 2.0f; P.radius[1] = 2.0f;
    for (int i = 0; i < 16; i++) P.weights[i] = 1.0f / (1 + i);
    P.color = color; P.normal = nrm; P.position = pos; P.out_f32 = nullptr; P.out_u8 = out;
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    for (int i = 0; i < 5; i++) sailLaunchPresent(P, pad, pad2, 0);
    cudaEventRecord(a);
    for (int i = 0; i < 50; i++) sailLaunchPresent(P, pad, pad2, 0);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    unsigned sum = 0; std::vector<unsigned char> o(n * 4); cudaMemcpy(o.data(), out, n * 4, cudaMemcpyDeviceToHost);
    for (auto v : o) sum = sum * 31u + v;
    printf("filter %d: %.4f ms per present (pad pass included), checksum %08x, %s\n", filter, ms / 50, sum, cudaGetErrorString(cudaGetLastError()));
    return 0;
}
