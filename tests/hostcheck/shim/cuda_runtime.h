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

struct float4 { float x, y, z, w; };
struct uint4 { unsigned x, y, z, w; };
struct dim3 { unsigned x = 1, y = 1, z = 1; };
typedef int cudaError_t;
typedef struct CUstream_st *cudaStream_t;
static const cudaError_t cudaSuccess = 0;

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

// declared so that the kernel-body templates parse; never called on the host
extern dim3 threadIdx, blockIdx, blockDim;
float __shfl_sync(unsigned, float, int);
unsigned __shfl_sync(unsigned, unsigned, int);
unsigned __reduce_add_sync(unsigned, unsigned);
unsigned __ballot_sync(unsigned, int);
int __popc(unsigned);
int __ffs(int);
unsigned long long atomicAdd(unsigned long long *, unsigned long long);
void __syncthreads();
