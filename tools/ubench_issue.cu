// ubench_issue.cu — dev tool (not part of the product library): measures warp-instruction issue rates
// on sm_100a for the instruction kinds the SDF interpreter is made of, to decide whether packing two
// rays into f32x2 operations (FFMA2/FADD2/FMUL2) and 3-input min/max (FMNMX3) removes issue slots.
// Prints warp-instructions per clock per SM (4 schedulers => 4.0 is the issue ceiling).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench_issue tools/ubench_issue.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstring>

constexpr int CHAINS = 8;     // independent dependency chains per thread
constexpr int ITERS = 4096;
constexpr int THREADS = 512;  // 16 warps per SM, 4 per scheduler

typedef unsigned long long u64;
__device__ __forceinline__ u64 pack(float a, float b) { return ((u64)__float_as_uint(b) << 32) | __float_as_uint(a); }

enum Kind { FFMA1, FADD1, FMUL1, FFMA2K, FADD2K, FMUL2K, MNMX, MNMX3, SQRT, MIX_ADD2_MNMX, MIX_ADD1_MNMX, MIX_FMA2_MNMX3, N_KIND };
static const char *kNames[N_KIND] = {"FFMA", "FADD", "FMUL", "FFMA2", "FADD2", "FMUL2", "FMNMX", "FMNMX3", "MUFU.SQRT",
                                     "FADD2+FMNMX (1:1)", "FADD+FMNMX (1:1)", "FFMA2+FMNMX3 (1:1)"};
// warp instructions issued per chain per iteration
static const int kInstPerStep[N_KIND] = {1, 1, 1, 1, 1, 1, 1, 1, 1, 2, 2, 2};

template <int K>
__global__ void __launch_bounds__(THREADS) k_issue(float seed, float *out, long long *cycles) {
    float a[CHAINS], m[CHAINS];
    u64 p[CHAINS];
    const float c0 = seed * 1.0001f, c1 = seed * 0.5f;
    const u64 pc0 = pack(c0, c1), pc1 = pack(c1, c0);
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) {
        a[i] = seed + threadIdx.x * 1e-3f + i;
        m[i] = a[i] * 0.25f;
        p[i] = pack(a[i], m[i]);
    }
    __syncthreads();
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < CHAINS; ++i) {
            if (K == FFMA1) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[i]) : "f"(c0), "f"(c1));
            if (K == FADD1) asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(c0));
            if (K == FMUL1) asm volatile("mul.rn.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(c0));
            if (K == FFMA2K) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(pc0), "l"(pc1));
            if (K == FADD2K) asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(p[i]) : "l"(pc0));
            if (K == FMUL2K) asm volatile("mul.rn.f32x2 %0, %0, %1;" : "+l"(p[i]) : "l"(pc0));
            if (K == MNMX) asm volatile("max.f32 %0, %0, %1;" : "+f"(m[i]) : "f"(c0));
            if (K == MNMX3) asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(m[i]) : "f"(c0), "f"(c1));
            if (K == SQRT) asm volatile("sqrt.approx.ftz.f32 %0, %0;" : "+f"(a[i]));
            if (K == MIX_ADD2_MNMX) {
                asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(p[i]) : "l"(pc0));
                asm volatile("max.f32 %0, %0, %1;" : "+f"(m[i]) : "f"(c0));
            }
            if (K == MIX_ADD1_MNMX) {
                asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(c0));
                asm volatile("max.f32 %0, %0, %1;" : "+f"(m[i]) : "f"(c0));
            }
            if (K == MIX_FMA2_MNMX3) {
                asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(pc0), "l"(pc1));
                asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(m[i]) : "f"(c0), "f"(c1));
            }
        }
    }
    const long long t1 = clock64();
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) s += a[i] + m[i] + __uint_as_float((unsigned)p[i]) + __uint_as_float((unsigned)(p[i] >> 32));
    out[blockIdx.x * THREADS + threadIdx.x] = s;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int K>
void run(int sms, float *d_out, long long *d_cyc) {
    long long h[1024];
    k_issue<K><<<sms, THREADS>>>(1.0f, d_out, d_cyc);  // warm
    k_issue<K><<<sms, THREADS>>>(1.0f, d_out, d_cyc);
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("%s: launch failed\n", kNames[K]); return; }
    cudaMemcpy(h, d_cyc, sms * sizeof(long long), cudaMemcpyDeviceToHost);
    double cyc = 0;
    for (int i = 0; i < sms; ++i) cyc += (double)h[i];
    cyc /= sms;
    const double warp_inst = (double)ITERS * CHAINS * kInstPerStep[K] * (THREADS / 32);
    printf("%-22s %8.3f warp-inst/clk/SM   (%.0f cycles per block)\n", kNames[K], warp_inst / cyc, cyc);
}

int main() {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    if (sms <= 0 || sms > 1024) { printf("no device\n"); return 1; }
    float *d_out; long long *d_cyc;
    cudaMalloc(&d_out, (size_t)sms * THREADS * sizeof(float));
    cudaMalloc(&d_cyc, sms * sizeof(long long));
    printf("SMs=%d, %d warps/SM, %d independent chains per thread; issue ceiling = 4.0\n", sms, THREADS / 32, CHAINS);
    run<FFMA1>(sms, d_out, d_cyc); run<FADD1>(sms, d_out, d_cyc); run<FMUL1>(sms, d_out, d_cyc);
    run<FFMA2K>(sms, d_out, d_cyc); run<FADD2K>(sms, d_out, d_cyc); run<FMUL2K>(sms, d_out, d_cyc);
    run<MNMX>(sms, d_out, d_cyc); run<MNMX3>(sms, d_out, d_cyc); run<SQRT>(sms, d_out, d_cyc);
    run<MIX_ADD2_MNMX>(sms, d_out, d_cyc); run<MIX_ADD1_MNMX>(sms, d_out, d_cyc); run<MIX_FMA2_MNMX3>(sms, d_out, d_cyc);
    return 0;
}
