// cuda_runtime.h — a stand-in that lets g++ compile the DEVICE headers of ray-marching_b200/csrc for the host.
// TEST INFRASTRUCTURE ONLY (tests/hostcheck): it exists so that the control flow of the decoded-program interpreters and
// of run-time generated map_scene code can be checked against the oracle where there is no GPU.  Nothing here is part
// of, linked into or loaded by the product library; the product has no CPU path.
// Arithmetic: the *_rn intrinsics are plain IEEE float operations (compile with -ffp-contract=off), so the EXACT mode
// of the device code runs bit for bit as on the GPU; fast-mode approximations (rsqrt, approximate sqrt and division)
// are replaced by their correctly rounded counterparts, so fast mode is only close, not identical.
#pragma once
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <cmath>
#include <cstring>
#include <string>
#include <vector>

#define RM_HOST_CHECK 1
#define __device__
#define __host__
#define __global__
#define __forceinline__ inline __attribute__((always_inline))
#define __noinline__ __attribute__((noinline))
#define __launch_bounds__(...)
#define __grid_constant__
#define __constant__ /* a plain global of the host build */

struct float4 { float x, y, z, w; };
struct uint4 { unsigned x, y, z, w; };
struct dim3 { unsigned x = 1, y = 1, z = 1; };
typedef int cudaError_t;
typedef struct CUstream_st *cudaStream_t;
static const cudaError_t cudaSuccess = 0;

static inline float __saturatef(float x) { return fminf(fmaxf(x, 0.0f), 1.0f); }  /* NaN -> 0 like the device's .sat */
static inline float __fadd_rn(float a, float b) { return a + b; }
static inline float __fsub_rn(float a, float b) { return a - b; }
static inline float __fmul_rn(float a, float b) { return a * b; }
static inline float __fdiv_rn(float a, float b) { return a / b; }
static inline float __fsqrt_rn(float a) { return sqrtf(a); }
static inline float __fdividef(float a, float b) { return a / b; }
static inline float rsqrtf(float a) { return 1.0f / sqrtf(a); }
static inline uint32_t __float_as_uint(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline float __uint_as_float(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static inline int __float2int_rn(float f) {  // cvt.rni.s32.f32: round to nearest even, saturating, NaN -> 0
    if (f != f) return 0;
    if (f >= 2147483648.0f) return INT32_MAX;
    if (f <= -2147483648.0f) return INT32_MIN;
    return (int)rintf(f);
}

// ---- one warp = 32 host threads in lock-step at the warp-collective calls --------------------------------------------
// tests/hostcheck/hostcheck.cpp runs a kernel body once per lane on its own thread (hc_run_grid); threadIdx / blockIdx
// are per-thread, and the warp collectives exchange values through a slot array guarded by a barrier.  All 32 lanes of
// a warp reach every collective (the kernel bodies only use them in warp-uniform control flow with a full mask).
#include <pthread.h>

struct HcWarp {
    pthread_barrier_t bar;
    unsigned slot[32];
};
struct dim3;
extern thread_local dim3 threadIdx, blockIdx, blockDim;  // defined (inline) in hc_grid.h
extern thread_local HcWarp *hc_warp;                     // the warp this lane belongs to
extern thread_local int hc_lane;

static inline unsigned hc_exchange(unsigned v, int src) {
    hc_warp->slot[hc_lane] = v;
    pthread_barrier_wait(&hc_warp->bar);
    const unsigned r = hc_warp->slot[src & 31];
    pthread_barrier_wait(&hc_warp->bar);
    return r;
}
static inline unsigned __shfl_sync(unsigned, unsigned v, int src) { return hc_exchange(v, src); }
static inline float __shfl_sync(unsigned, float v, int src) { return __uint_as_float(hc_exchange(__float_as_uint(v), src)); }
static inline float __shfl_xor_sync(unsigned, float v, int lane_mask) { return __uint_as_float(hc_exchange(__float_as_uint(v), hc_lane ^ lane_mask)); }
static inline unsigned __reduce_add_sync(unsigned, unsigned v) {
    hc_warp->slot[hc_lane] = v;
    pthread_barrier_wait(&hc_warp->bar);
    unsigned s = 0;
    for (int i = 0; i < 32; ++i) s += hc_warp->slot[i];
    pthread_barrier_wait(&hc_warp->bar);
    return s;
}
static inline unsigned __ballot_sync(unsigned, int pred) {
    hc_warp->slot[hc_lane] = pred ? 1u : 0u;
    pthread_barrier_wait(&hc_warp->bar);
    unsigned m = 0;
    for (int i = 0; i < 32; ++i) m |= hc_warp->slot[i] << i;
    pthread_barrier_wait(&hc_warp->bar);
    return m;
}
static inline int __popc(unsigned x) { return __builtin_popcount(x); }
static inline int __ffs(int x) { return __builtin_ffs(x); }
static inline unsigned long long atomicAdd(unsigned long long *p, unsigned long long v) { return __atomic_fetch_add(p, v, __ATOMIC_RELAXED); }
void __syncthreads();  // block-wide barriers only occur in the kernel wrappers (.cu files), which the host check does not compile
