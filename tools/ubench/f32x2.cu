// Micro-benchmark: does the packed FP32 pipe (FMUL2 / FADD2 / FFMA2, sm_100) free issue slots?
// Each kernel runs ITER iterations of 8 independent dependency chains per thread.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -fmad=false -o f32x2 f32x2.cu && ./f32x2
#include <cstdio>
#include <cuda_runtime.h>
#define ITER 4096
#define CH 8

typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b) { u64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void upk(u64 v, float& a, float& b) { asm("mov.b64 {%0,%1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ u64 mul2(u64 a, u64 b) { u64 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
// a + b as fma(a, ONE, b): ptxas contracts mul.rn.f32x2 + add.rn.f32x2 into one FFMA2 (single rounding) even with
// -fmad=false; with ONE a run-time value it cannot, and a*1+b rounds exactly like a+b.
__device__ __forceinline__ u64 add2(u64 a, u64 b, u64 one) { u64 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(one), "l"(b)); return r; }

template <int MODE>
__global__ void __launch_bounds__(128, 8) k(float* out, float seed, int salt) {
    const u64 one = pk(seed, seed);  // seed = 1.0f at run time
    float a[CH], b[CH];
    int c[CH];
    for (int i = 0; i < CH; i++) { a[i] = seed + i + threadIdx.x; b[i] = seed * 0.5f + i; c[i] = salt + i; }
    const float m = 0.999f + seed * 1e-9f, ad = seed * 1e-7f;
    for (int it = 0; it < ITER; it++) {
        if (MODE == 0) {  // scalar: 2 FMUL + 2 FADD per chain
#pragma unroll
            for (int i = 0; i < CH; i++) { a[i] = __fmul_rn(a[i], m); b[i] = __fmul_rn(b[i], m); a[i] = __fadd_rn(a[i], ad); b[i] = __fadd_rn(b[i], ad); }
        } else if (MODE == 1) {  // packed: 1 FMUL2 + 1 FADD2 per chain
#pragma unroll
            for (int i = 0; i < CH; i++) {
                u64 v = add2(mul2(pk(a[i], b[i]), pk(m, m)), pk(ad, ad), one);
                upk(v, a[i], b[i]);
            }
        } else if (MODE == 2) {  // scalar FP + integer ALU work: 2 FMUL + 2 FADD + 4 integer ops
#pragma unroll
            for (int i = 0; i < CH; i++) {
                a[i] = __fmul_rn(a[i], m); b[i] = __fmul_rn(b[i], m); a[i] = __fadd_rn(a[i], ad); b[i] = __fadd_rn(b[i], ad);
                c[i] = (c[i] ^ (c[i] << 3)) + salt; c[i] = (c[i] ^ (c[i] >> 5)) + it;
            }
        } else {  // packed FP + the same integer work
#pragma unroll
            for (int i = 0; i < CH; i++) {
                u64 v = add2(mul2(pk(a[i], b[i]), pk(m, m)), pk(ad, ad), one);
                upk(v, a[i], b[i]);
                c[i] = (c[i] ^ (c[i] << 3)) + salt; c[i] = (c[i] ^ (c[i] >> 5)) + it;
            }
        }
    }
    float s = 0;
    for (int i = 0; i < CH; i++) s += a[i] + b[i] + (float)c[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE>
float run(float* d, int grid) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<grid, 128>>>(d, 1.0f, 3);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    for (int r = 0; r < 5; r++) k<MODE><<<grid, 128>>>(d, 1.0f, 3);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    return ms / 5;
}

int main() {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    int grid = p.multiProcessorCount * 8;
    float* d;
    cudaMalloc(&d, (size_t)grid * 128 * sizeof(float));
    float t0 = run<0>(d, grid), t1 = run<1>(d, grid), t2 = run<2>(d, grid), t3 = run<3>(d, grid);
    double fp = (double)grid * 128 * ITER * CH * 4;  // FP32 lane-ops per launch
    printf("SMs %d\n", p.multiProcessorCount);
    printf("scalar FMUL/FADD        : %.3f ms  %.2f Tops/s\n", t0, fp / t0 / 1e9);
    printf("packed FMUL2/FADD2      : %.3f ms  %.2f Tops/s\n", t1, fp / t1 / 1e9);
    printf("scalar FP + integer ALU : %.3f ms  %.2f Tops/s (FP part)\n", t2, fp / t2 / 1e9);
    printf("packed FP + integer ALU : %.3f ms  %.2f Tops/s (FP part)\n", t3, fp / t3 / 1e9);
    return 0;
}
