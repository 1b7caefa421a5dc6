// ubench_ffma2.cu — measurement only: FP32 FMA throughput of one B200 with scalar FFMA and with the packed FFMA2
// (fma.rn.f32x2), same harness for both: 16 independent accumulator chains per thread, 148 x 8 blocks of 256 threads.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_ffma2 tools/ubench_ffma2.cu && ./ubench_ffma2
// Prints TFLOP/s (2 flops per lane-FMA) for both, and the ratio.  Question it answers: does a packed instruction double the
// FP32 rate (then the FFMA probe behind roofline.peak would understate the peak) or is it one issue slot for two passes
// of the same pipe (then the peak stands and only issue slots are saved)?
#include <cstdio>
#include <cuda_runtime.h>

typedef unsigned long long u64;
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 r; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ u64 pk(float a, float b) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ float lo(u64 v) { float a, b; asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); return a + b; }

constexpr int CHAINS = 16, ITERS = 4096;

__global__ void k_scalar(float *out, float x, float y) {
    float acc[CHAINS];
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) acc[c] = threadIdx.x * 1e-3f + c;
    for (int i = 0; i < ITERS; ++i) {
#pragma unroll
        for (int c = 0; c < CHAINS; ++c) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(acc[c]) : "f"(x), "f"(y));
    }
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) s += acc[c];
    if (s == 12345.678f) out[0] = s;
}

__global__ void k_packed(float *out, float x, float y) {
    u64 acc[CHAINS];
    const u64 X = pk(x, x), Y = pk(y, y);
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) acc[c] = pk(threadIdx.x * 1e-3f + c, c * 0.5f);
    for (int i = 0; i < ITERS; ++i) {
#pragma unroll
        for (int c = 0; c < CHAINS; ++c) acc[c] = fma2(acc[c], X, Y);
    }
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) s += lo(acc[c]);
    if (s == 12345.678f) out[0] = s;
}

int main() {
    float *out;
    cudaMalloc(&out, 4);
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    const int blocks = 148 * 8, threads = 256;
    double tf[2] = {0, 0};
    for (int which = 0; which < 2; ++which) {
        float best = 1e30f;
        for (int rep = 0; rep < 6; ++rep) {
            cudaEventRecord(a);
            if (which == 0) k_scalar<<<blocks, threads>>>(out, 0.999f, 1e-4f);
            else k_packed<<<blocks, threads>>>(out, 0.999f, 1e-4f);
            cudaEventRecord(b);
            cudaEventSynchronize(b);
            float ms = 0;
            cudaEventElapsedTime(&ms, a, b);
            if (rep >= 2 && ms < best) best = ms;
        }
        const double lane_fmas = (double)blocks * threads * CHAINS * ITERS * (which ? 2.0 : 1.0);
        tf[which] = 2.0 * lane_fmas / (best * 1e-3) / 1e12;
        printf("%s: %.3f ms, %.2f TFLOP/s\n", which ? "FFMA2 (fma.rn.f32x2)" : "FFMA  (fma.rn.f32)  ", best, tf[which]);
    }
    printf("ratio packed / scalar: %.3f\n", tf[1] / tf[0]);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { printf("CUDA error: %s\n", cudaGetErrorString(e)); return 1; }
    return 0;
}
