// ubench_sdf.cu — dev tool (not part of the product library): the interpreter's real per-instruction
// sequences (sphere+smooth-union, box+smooth-union) for R=8 rays per lane, written three ways:
//   SCALAR   what rm_device.cuh's Arith<false> compiles to today (FADD/FMUL/FFMA/FMNMX/MUFU)
//   MNMX3    the same with 3-input min/max where two min/max ops are adjacent
//   PACKED   rays paired into f32x2 registers: FADD2/FMUL2/FFMA2 for sub/mul/fma, scalar MUFU and min/max
// The instruction operands come from shared memory (one LDS.128 x3 per instruction, as in the kernel).
// Prints cycles per (warp x fused instruction x 8 rays); lower is better.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench_sdf tools/ubench_sdf.cu
#include <cuda_runtime.h>
#include <cstdio>

constexpr int R = 8, NI = 32, REPS = 4096, THREADS = 128, CTAS_PER_SM = 4;
typedef unsigned long long u64;

__device__ __forceinline__ float fsqrt(float a) { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a)); return r; }
__device__ __forceinline__ float max3(float a, float b, float c) { float r; asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r; }
__device__ __forceinline__ u64 pk(float lo, float hi) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void upk(u64 v, float &lo, float &hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ u64 add2(u64 a, u64 b) { u64 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ u64 mul2(u64 a, u64 b) { u64 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }

// ---- scalar forms (Arith<false>) ------------------------------------------------------------
__device__ __forceinline__ float smooth1(float a, float b, float k, float ik, float kq) {
    float e = fmaxf(k - fabsf(a - b), 0.f), h = e * ik;
    return fmaf(-(h * h), kq, fminf(a, b));
}
__device__ __forceinline__ float sphere1(float x, float y, float z, float4 q0) {
    float dx = x - q0.x, dy = y - q0.y, dz = z - q0.z;
    return fsqrt(fmaf(dz, dz, fmaf(dy, dy, dx * dx))) - q0.w;
}
template <bool M3>
__device__ __forceinline__ float box1(float x, float y, float z, float4 q0, float4 q1) {
    float qx = fabsf(x - q0.x) - q0.w, qy = fabsf(y - q0.y) - q1.x, qz = fabsf(z - q0.z) - q1.y;
    float mx = fmaxf(qx, 0.f), my = fmaxf(qy, 0.f), mz = fmaxf(qz, 0.f);
    float len = fsqrt(fmaf(mz, mz, fmaf(my, my, mx * mx)));
    float in = M3 ? fminf(max3(qx, qy, qz), 0.f) : fminf(fmaxf(qx, fmaxf(qy, qz)), 0.f);
    return len + in;
}

// ---- packed forms: two rays per 64-bit register ------------------------------------------------
// negated centre / radius are pre-broadcast per instruction (2 MOV-pairs amortised over 4 ray pairs)
__device__ __forceinline__ u64 sphere2(u64 x, u64 y, u64 z, u64 ncx, u64 ncy, u64 ncz, float r) {
    u64 dx = add2(x, ncx), dy = add2(y, ncy), dz = add2(z, ncz);
    u64 s = fma2(dz, dz, fma2(dy, dy, mul2(dx, dx)));
    float s0, s1; upk(s, s0, s1);
    return pk(fsqrt(s0) - r, fsqrt(s1) - r);  // the final subtract stays scalar (could be add2 too)
}
__device__ __forceinline__ u64 smooth2(u64 a, u64 b, float k, u64 ik2, u64 nkq2) {
    float a0, a1, b0, b1; upk(a, a0, a1); upk(b, b0, b1);
    float e0 = fmaxf(k - fabsf(a0 - b0), 0.f), e1 = fmaxf(k - fabsf(a1 - b1), 0.f);  // |.| modifier: scalar only
    u64 h = mul2(pk(e0, e1), ik2);
    return fma2(mul2(h, h), nkq2, pk(fminf(a0, b0), fminf(a1, b1)));
}
__device__ __forceinline__ u64 box2(u64 x, u64 y, u64 z, u64 ncx, u64 ncy, u64 ncz, float4 q0, float4 q1) {
    float dx0, dx1, dy0, dy1, dz0, dz1;
    upk(add2(x, ncx), dx0, dx1); upk(add2(y, ncy), dy0, dy1); upk(add2(z, ncz), dz0, dz1);
    float qx0 = fabsf(dx0) - q0.w, qx1 = fabsf(dx1) - q0.w, qy0 = fabsf(dy0) - q1.x, qy1 = fabsf(dy1) - q1.x,
          qz0 = fabsf(dz0) - q1.y, qz1 = fabsf(dz1) - q1.y;
    u64 mx = pk(fmaxf(qx0, 0.f), fmaxf(qx1, 0.f)), my = pk(fmaxf(qy0, 0.f), fmaxf(qy1, 0.f)), mz = pk(fmaxf(qz0, 0.f), fmaxf(qz1, 0.f));
    float s0, s1; upk(fma2(mz, mz, fma2(my, my, mul2(mx, mx))), s0, s1);
    return add2(pk(fsqrt(s0), fsqrt(s1)), pk(fminf(max3(qx0, qy0, qz0), 0.f), fminf(max3(qx1, qy1, qz1), 0.f)));
}

enum Variant { SCALAR, MNMX3V, PACKED };
template <int V, bool BOX>
__global__ void __launch_bounds__(THREADS, CTAS_PER_SM) k_sdf(const float4 *prog_g, float *out, long long *cycles) {
    __shared__ float4 prog[3 * NI];
    for (int i = threadIdx.x; i < 3 * NI; i += THREADS) prog[i] = prog_g[i];
    float px[R], py[R], pz[R], acc[R];
#pragma unroll
    for (int r = 0; r < R; ++r) { px[r] = 0.01f * threadIdx.x + r; py[r] = 0.3f * r - 1.f; pz[r] = 0.02f * threadIdx.x; acc[r] = 1e3f; }
    u64 X[R / 2], Y[R / 2], Z[R / 2], A[R / 2];
#pragma unroll
    for (int r = 0; r < R / 2; ++r) { X[r] = pk(px[2 * r], px[2 * r + 1]); Y[r] = pk(py[2 * r], py[2 * r + 1]); Z[r] = pk(pz[2 * r], pz[2 * r + 1]); A[r] = pk(1e3f, 1e3f); }
    __syncthreads();
    const long long t0 = clock64();
#pragma unroll 1
    for (int rep = 0; rep < REPS; ++rep) {
#pragma unroll 1
        for (int i = 0; i < NI; ++i) {
            const float4 q0 = prog[3 * i], q1 = prog[3 * i + 1], q2 = prog[3 * i + 2];
            if (V == PACKED) {
                const u64 ncx = pk(-q0.x, -q0.x), ncy = pk(-q0.y, -q0.y), ncz = pk(-q0.z, -q0.z), ik2 = pk(q1.w, q1.w), nkq2 = pk(-q2.x, -q2.x);
#pragma unroll
                for (int r = 0; r < R / 2; ++r) {
                    u64 v = BOX ? box2(X[r], Y[r], Z[r], ncx, ncy, ncz, q0, q1) : sphere2(X[r], Y[r], Z[r], ncx, ncy, ncz, q0.w);
                    A[r] = smooth2(A[r], v, q1.z, ik2, nkq2);
                }
            } else {
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    float v = BOX ? box1<V == MNMX3V>(px[r], py[r], pz[r], q0, q1) : sphere1(px[r], py[r], pz[r], q0);
                    acc[r] = smooth1(acc[r], v, q1.z, q1.w, q2.x);
                }
            }
        }
    }
    const long long t1 = clock64();
    float s = 0.f;
#pragma unroll
    for (int r = 0; r < R; ++r) s += acc[r];
#pragma unroll
    for (int r = 0; r < R / 2; ++r) { float a, b; upk(A[r], a, b); s += a + b; }
    out[blockIdx.x * THREADS + threadIdx.x] = s;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int V, bool BOX>
void run(const char *name, int sms, const float4 *d_prog, float *d_out, long long *d_cyc) {
    const int grid = sms * CTAS_PER_SM;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k_sdf<V, BOX><<<grid, THREADS>>>(d_prog, d_out, d_cyc);  // warm-up
    float best = 1e30f, worst = 0.f;
    for (int t = 0; t < 3; ++t) {
        cudaEventRecord(e0);
        k_sdf<V, BOX><<<grid, THREADS>>>(d_prog, d_out, d_cyc);
        cudaEventRecord(e1);
        if (cudaEventSynchronize(e1) != cudaSuccess) { printf("%s: failed\n", name); return; }
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        best = ms < best ? ms : best; worst = ms > worst ? ms : worst;
    }
    // whole-kernel time per (warp x fused instruction of 8 rays), all SMs: ns
    const double units = (double)grid * (THREADS / 32) * REPS * NI;
    printf("%-28s %8.3f ms (worst %8.3f)   %7.2f ps per (warp, fused instruction, 8 rays)\n", name, best, worst, best * 1e9 / units);
}

int main() {
    int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    if (sms <= 0) { printf("no device\n"); return 1; }
    float4 hp[3 * NI];
    for (int i = 0; i < NI; ++i) { hp[3 * i] = make_float4(0.1f * i - 1.f, 0.05f * i, 0.3f, 0.4f); hp[3 * i + 1] = make_float4(0.3f, 0.35f, 0.25f, 4.f); hp[3 * i + 2] = make_float4(0.0625f, 0, 0, 0); }
    float4 *d_prog; float *d_out; long long *d_cyc;
    cudaMalloc(&d_prog, sizeof(hp)); cudaMemcpy(d_prog, hp, sizeof(hp), cudaMemcpyHostToDevice);
    cudaMalloc(&d_out, (size_t)sms * CTAS_PER_SM * THREADS * sizeof(float)); cudaMalloc(&d_cyc, sms * CTAS_PER_SM * sizeof(long long));
    printf("SMs=%d, %d CTAs/SM x %d threads (one warp of each CTA per scheduler), R=%d\n", sms, CTAS_PER_SM, THREADS, R);
    run<SCALAR, false>("sphere+smooth scalar", sms, d_prog, d_out, d_cyc);
    run<PACKED, false>("sphere+smooth packed f32x2", sms, d_prog, d_out, d_cyc);
    run<SCALAR, true>("box+smooth scalar", sms, d_prog, d_out, d_cyc);
    run<MNMX3V, true>("box+smooth scalar + max3", sms, d_prog, d_out, d_cyc);
    run<PACKED, true>("box+smooth packed + max3", sms, d_prog, d_out, d_cyc);
    return 0;
}
