// rm_device.cuh — device-side building blocks shared by all kernels of the render pass.
// Arithmetic restates /root/reference/src/ray_marching/ray_marching.wgsl; each function cites
// the shader lines it follows.  Two arithmetic policies:
//   Arith<true>  (RM_MODE_EXACT): every add/mul/div/sqrt rounds once (no FMA contraction, IEEE
//                sqrt.rn/div.rn) in the shader's evaluation order -> bit-identical to the oracle.
//   Arith<false> (RM_MODE_FAST):  FMA contraction, MUFU sqrt/rsqrt/rcp.
#pragma once
#ifndef __CUDACC_RTC__
#include <cuda_runtime.h>
#include <stdint.h>
#endif

#include "rm_program.h"

#define RM_FOR_R _Pragma("unroll") for (int r = 0; r < R; ++r)
#define RM_FOR_H _Pragma("unroll") for (int h = 0; h < R / 2; ++h)  /* pairs of rays (f2_t) */

// The few places that pin an instruction with inline PTX go through these macros.  RM_HOST_CHECK is defined only by
// tests/hostcheck (a host compile of the interpreters and of generated map_scene code, checked against the oracle
// without a GPU); the library itself is never built with it.
#ifndef RM_HOST_CHECK
#define RM_PTX_SQRT_APPROX(r, a) asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a))
#define RM_PTX_MIN_ACC(acc, v) asm("min.f32 %0, %0, %1;" : "+f"(acc) : "f"(v))
#define RM_PTX_MAX_ACC(acc, v) asm("max.f32 %0, %0, %1;" : "+f"(acc) : "f"(v))
#define RM_PTX_SUB_ACC(acc, v) asm("sub.rn.f32 %0, %0, %1;" : "+f"(acc) : "f"(v))
#define RM_PTX_FMA_ACC(acc, a, b) asm("fma.rn.f32 %0, %1, %2, %0;" : "+f"(acc) : "f"(a), "f"(b))
#define RM_OPAQUE_F3(a, b, c) asm volatile("" : "+f"(a), "+f"(b), "+f"(c))  /* scheduling fence: values become opaque */
/* the block's tile coordinates, read again where they are needed so that nothing derived from them lives across the loops */
#define RM_REREAD_CTAID(cx_, cy_) asm volatile("mov.u32 %0, %%ctaid.x;\n\tmov.u32 %1, %%ctaid.y;" : "=r"(cx_), "=r"(cy_))
#else
#define RM_PTX_SQRT_APPROX(r, a) ((r) = sqrtf(a))
#define RM_PTX_MIN_ACC(acc, v) ((acc) = fminf((acc), (v)))
#define RM_PTX_MAX_ACC(acc, v) ((acc) = fmaxf((acc), (v)))
#define RM_PTX_SUB_ACC(acc, v) ((acc) = (acc) - (v))
#define RM_PTX_FMA_ACC(acc, a, b) ((acc) = fmaf((a), (b), (acc)))
#define RM_OPAQUE_F3(a, b, c) ((void)0)
#define RM_REREAD_CTAID(cx_, cy_) ((cx_) = blockIdx.x, (cy_) = blockIdx.y)
#endif

namespace rm {

// constants of the shader (SURVEY.md §5 "Config / flags")
#define RM_NORMAL_EPS 0.0001f                   /* wgsl:136 */
#define RM_LIGHT_X 2.0f                         /* wgsl:100 */
#define RM_LIGHT_Y (-5.0f)
#define RM_LIGHT_Z 3.0f
#define RM_AMBIENT 0.02f                        /* wgsl:103 */
#define RM_ALBEDO_R 0.4f                        /* wgsl:105 */
#define RM_ALBEDO_G 0.7f
#define RM_ALBEDO_B 0.1f
#define RM_FLOOR_Y (-1.5f)                      /* wgsl:119 */

// Per-launch frame block: the reference's two uniform buffers + the band being shaded.
struct FrameParams {
    float inv_proj[16];  // column-major, Uniforms @16   wgsl:26
    float inv_view[16];  // column-major, Uniforms @80   wgsl:28
    float extent[2];     // Uniforms @0                  wgsl:24
    float min_dist, max_dist;  // RayMarchLimits         wgsl:78-82
    uint32_t max_iter;
    uint32_t W, H, row0, nrows;
    // interleaved strips (multi-GPU load balance): local row ry of the band maps to frame row
    // row0 + (ry / strip_rows) * strip_pitch + ry % strip_rows; strip_pitch == 0 = contiguous band
    uint32_t strip_rows, strip_pitch;
    uint32_t n_insn;     // decoded instruction count (0 = empty scene)
    // 0: dst holds the band's rows packed (local row ry -> dst row ry).  1: dst is the WHOLE frame and every row lands
    // at its frame position (frame_row(ry)) — used to store strips straight into another GPU's frame over NVLink.
    uint32_t dst_frame_rows;
    // Per-frame constants of ray generation, computed once on the host with the same IEEE float operations the
    // shader's expressions denote (rm_abi.cu make_params), so they are bit-identical to evaluating them per ray:
    float ro4[4];        // inv_view * (0,0,0,1)                        wgsl:39-40
    float sample_ax[4];  // ((i + 0.5)/4 - 0.5) / extent.x * 2, i = 0..3   wgsl:52-53
    float sample_ay[4];  // ((j + 0.5)/4 - 0.5) / extent.y * 2, j = 0..3
    // FAST mode only: wgsl:52-61 is affine in the sample position, so the un-normalised direction of sample (i, j) of pixel
    // (px, py) is  ray_c0 + X * ray_cx + Y * ray_cy  with  X = px + ray_bx + i * ray_gx,  Y = py + ray_by + j * ray_gy
    // (for extent == (W, H): X = px + (2i+1)/8).  The three 4-vectors are inv_view * inv_proj composed in double on the host
    // (rm_frame.h): 8 FFMA per sample instead of two 4x4 matrix-vector products and the pixel -> pt_screen divisions.
    float ray_c0[4], ray_cx[4], ray_cy[4];
    float ray_gx, ray_gy, ray_bx, ray_by;
};

__device__ __forceinline__ uint32_t frame_row(const FrameParams &fp, uint32_t ry) {
    if (fp.strip_pitch == 0) return fp.row0 + ry;
    return fp.row0 + (ry / fp.strip_rows) * fp.strip_pitch + ry % fp.strip_rows;
}

// row of dst that local row ry of the band is stored to
__device__ __forceinline__ uint32_t dst_row(const FrameParams &fp, uint32_t ry) {
    return fp.dst_frame_rows ? frame_row(fp, ry) : ry;
}

template <bool EX>
struct Arith;

template <>
struct Arith<true> {
    static __device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
    static __device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
    static __device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
    static __device__ __forceinline__ float div(float a, float b) { return __fdiv_rn(a, b); }
    static __device__ __forceinline__ float sqrt(float a) { return __fsqrt_rn(a); }
    // a*b + c with two roundings
    static __device__ __forceinline__ float mad(float a, float b, float c) { return __fadd_rn(__fmul_rn(a, b), c); }
    // (x*x + y*y) + z*z
    static __device__ __forceinline__ float dot3s(float x, float y, float z) {
        return __fadd_rn(__fadd_rn(__fmul_rn(x, x), __fmul_rn(y, y)), __fmul_rn(z, z));
    }
    static __device__ __forceinline__ float length3(float x, float y, float z) { return __fsqrt_rn(dot3s(x, y, z)); }
    static __device__ __forceinline__ void normalize3(float &x, float &y, float &z) {
        float l = length3(x, y, z);
        x = __fdiv_rn(x, l); y = __fdiv_rn(y, l); z = __fdiv_rn(z, l);
    }
};

template <>
struct Arith<false> {
    static __device__ __forceinline__ float add(float a, float b) { return a + b; }
    static __device__ __forceinline__ float sub(float a, float b) { return a - b; }
    static __device__ __forceinline__ float mul(float a, float b) { return a * b; }
    static __device__ __forceinline__ float div(float a, float b) { return __fdividef(a, b); }
    static __device__ __forceinline__ float sqrt(float a) {
        float r;
        RM_PTX_SQRT_APPROX(r, a);
        return r;
    }
    static __device__ __forceinline__ float mad(float a, float b, float c) { return fmaf(a, b, c); }
    static __device__ __forceinline__ float dot3s(float x, float y, float z) { return fmaf(z, z, fmaf(y, y, x * x)); }
    static __device__ __forceinline__ float length3(float x, float y, float z) { return sqrt(dot3s(x, y, z)); }
    static __device__ __forceinline__ void normalize3(float &x, float &y, float &z) {
        float r = rsqrtf(dot3s(x, y, z));
        x *= r; y *= r; z *= r;
    }
};

// ---- primitives -----------------------------------------------------------------------------
// wgsl:229-233  length(pos - center) - radius
template <bool EX>
__device__ __forceinline__ float sd_sphere(float px, float py, float pz, float cx, float cy, float cz, float r) {
    using A = Arith<EX>;
    float dx = A::sub(px, cx), dy = A::sub(py, cy), dz = A::sub(pz, cz);
    return A::sub(A::length3(dx, dy, dz), r);
}
// wgsl:235-240  q = abs(pos-c) - r;  length(max(q,0)) + min(max(q.x,max(q.y,q.z)),0)
template <bool EX>
__device__ __forceinline__ float sd_box(float px, float py, float pz, float cx, float cy, float cz, float rx,
                                        float ry, float rz) {
    using A = Arith<EX>;
    float qx = A::sub(fabsf(A::sub(px, cx)), rx);
    float qy = A::sub(fabsf(A::sub(py, cy)), ry);
    float qz = A::sub(fabsf(A::sub(pz, cz)), rz);
    float len = A::length3(fmaxf(qx, 0.0f), fmaxf(qy, 0.0f), fmaxf(qz, 0.0f));
    return A::add(len, fminf(fmaxf(qx, fmaxf(qy, qz)), 0.0f));
}

// ---- two rays at a time: the packed FP32 instructions of sm_100 (FADD2 / FMUL2 / FFMA2) -----------------------------
// add / mul / fma.rn.f32x2 work on a 64-bit register pair holding two floats and issue ONCE for both: the kernels here are
// bound by instruction issue, with the FMA pipe a third idle, so the generated scene code evaluates the rays of a thread
// in pairs.  ptxas folds a broadcast scalar (immediate or uniform register), |x| and -x of the halves into the packed
// instruction's operands (FADD2 R, |R|.F32x2.HI_LO, -UR.F32), so constants and modifiers cost nothing extra.  Per lane the
// operations, their order and their roundings are the scalar fast-mode ones (sd_sphere<false>, sd_box<false>,
// op_smooth<false, SAT>): the same bits.  Fast mode only; min / max / sqrt stay scalar per half (no packed form).
#ifndef RM_HOST_CHECK
typedef unsigned long long f2_t;
__device__ __forceinline__ f2_t f2_pack(float lo, float hi) { f2_t r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ float f2_lo(f2_t v) { float a, b; asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); return a; }
__device__ __forceinline__ float f2_hi(f2_t v) { float a, b; asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); return b; }
__device__ __forceinline__ f2_t f2_add(f2_t a, f2_t b) { f2_t r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f2_t f2_sub(f2_t a, f2_t b) { f2_t r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f2_t f2_mul(f2_t a, f2_t b) { f2_t r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f2_t f2_fma(f2_t a, f2_t b, f2_t c) { f2_t r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
#else
struct f2_t { float lo, hi; };
static inline f2_t f2_pack(float lo, float hi) { f2_t r; r.lo = lo; r.hi = hi; return r; }
static inline float f2_lo(f2_t v) { return v.lo; }
static inline float f2_hi(f2_t v) { return v.hi; }
static inline f2_t f2_add(f2_t a, f2_t b) { return f2_pack(a.lo + b.lo, a.hi + b.hi); }
static inline f2_t f2_sub(f2_t a, f2_t b) { return f2_pack(a.lo - b.lo, a.hi - b.hi); }
static inline f2_t f2_mul(f2_t a, f2_t b) { return f2_pack(a.lo * b.lo, a.hi * b.hi); }
static inline f2_t f2_fma(f2_t a, f2_t b, f2_t c) { return f2_pack(fmaf(a.lo, b.lo, c.lo), fmaf(a.hi, b.hi, c.hi)); }
#endif
__device__ __forceinline__ f2_t f2_all(float c) { return f2_pack(c, c); }
__device__ __forceinline__ f2_t f2_abs(f2_t v) { return f2_pack(fabsf(f2_lo(v)), fabsf(f2_hi(v))); }
__device__ __forceinline__ f2_t f2_neg(f2_t v) { return f2_pack(-f2_lo(v), -f2_hi(v)); }
__device__ __forceinline__ f2_t f2_min(f2_t a, f2_t b) { return f2_pack(fminf(f2_lo(a), f2_lo(b)), fminf(f2_hi(a), f2_hi(b))); }
__device__ __forceinline__ f2_t f2_max(f2_t a, f2_t b) { return f2_pack(fmaxf(f2_lo(a), f2_lo(b)), fmaxf(f2_hi(a), f2_hi(b))); }
__device__ __forceinline__ f2_t f2_sqrt(f2_t v) { return f2_pack(Arith<false>::sqrt(f2_lo(v)), Arith<false>::sqrt(f2_hi(v))); }
// sd_sphere<false> for two rays
__device__ __forceinline__ f2_t sd_sphere2(f2_t px, f2_t py, f2_t pz, float cx, float cy, float cz, float r) {
    const f2_t dx = f2_sub(px, f2_all(cx)), dy = f2_sub(py, f2_all(cy)), dz = f2_sub(pz, f2_all(cz));
    return f2_sub(f2_sqrt(f2_fma(dz, dz, f2_fma(dy, dy, f2_mul(dx, dx)))), f2_all(r));
}
// sd_box<false> for two rays
__device__ __forceinline__ f2_t sd_box2(f2_t px, f2_t py, f2_t pz, float cx, float cy, float cz, float rx, float ry, float rz) {
    const f2_t qx = f2_sub(f2_abs(f2_sub(px, f2_all(cx))), f2_all(rx));
    const f2_t qy = f2_sub(f2_abs(f2_sub(py, f2_all(cy))), f2_all(ry));
    const f2_t qz = f2_sub(f2_abs(f2_sub(pz, f2_all(cz))), f2_all(rz));
    const f2_t zero = f2_all(0.0f);
    const f2_t ux = f2_max(qx, zero), uy = f2_max(qy, zero), uz = f2_max(qz, zero);
    const f2_t len = f2_sqrt(f2_fma(uz, uz, f2_fma(uy, uy, f2_mul(ux, ux))));
    return f2_add(len, f2_min(f2_max(qx, f2_max(qy, qz)), zero));
}
// op_smooth<false, SAT> for two rays: (ks, s) = (k*s, s) of smooth_constants_fast
template <bool SAT>
__device__ __forceinline__ f2_t op_smooth2(f2_t a, f2_t b, float ks, float s) {
    const f2_t d = f2_sub(a, b);
    const float x0 = fmaf(-fabsf(f2_lo(d)), s, ks), x1 = fmaf(-fabsf(f2_hi(d)), s, ks);
    const f2_t e = SAT ? f2_pack(__saturatef(x0), __saturatef(x1)) : f2_pack(fmaxf(x0, 0.0f), fmaxf(x1, 0.0f));
    return f2_fma(f2_neg(e), e, f2_min(a, b));
}

// ---- N4 extension (include/rm_b200.h RM_OP_PLANE, RM_OP_*_PUSH; normative order: oracle/rm_oracle.c map_scene) ----
// plane: ((x*nx + y*ny) + z*nz) + h
template <bool EX>
__device__ __forceinline__ float sd_plane(float px, float py, float pz, float nx, float ny, float nz, float h) {
    using A = Arith<EX>;
    if (EX) return A::add(A::add(A::add(A::mul(px, nx), A::mul(py, ny)), A::mul(pz, nz)), h);
    return fmaf(pz, nz, fmaf(py, ny, fmaf(px, nx, h)));
}
// TranslationPush: q = b - t
template <bool EX>
__device__ __forceinline__ void xf_translate(float bx, float by, float bz, float tx, float ty, float tz, float &qx, float &qy,
                                             float &qz) {
    using A = Arith<EX>;
    qx = A::sub(bx, tx); qy = A::sub(by, ty); qz = A::sub(bz, tz);
}
// RotationPush: q = transpose(M) * b, M row-major: qx = (m00*bx + m10*by) + m20*bz ...
template <bool EX>
__device__ __forceinline__ void xf_rotate(float bx, float by, float bz, float m00, float m01, float m02, float m10, float m11,
                                          float m12, float m20, float m21, float m22, float &qx, float &qy, float &qz) {
    using A = Arith<EX>;
    if (EX) {
        const float x = A::add(A::add(A::mul(m00, bx), A::mul(m10, by)), A::mul(m20, bz));
        const float y = A::add(A::add(A::mul(m01, bx), A::mul(m11, by)), A::mul(m21, bz));
        const float z = A::add(A::add(A::mul(m02, bx), A::mul(m12, by)), A::mul(m22, bz));
        qx = x; qy = y; qz = z;
    } else {
        const float x = fmaf(m20, bz, fmaf(m10, by, m00 * bx));
        const float y = fmaf(m21, bz, fmaf(m11, by, m01 * bx));
        const float z = fmaf(m22, bz, fmaf(m12, by, m02 * bx));
        qx = x; qy = y; qz = z;
    }
}
// ScalePush: q = b / s (fast mode: b * (1/s), the reciprocal computed once per program on the host)
template <bool EX>
__device__ __forceinline__ void xf_scale(float bx, float by, float bz, float s, float inv_s, float &qx, float &qy, float &qz) {
    if (EX) { qx = __fdiv_rn(bx, s); qy = __fdiv_rn(by, s); qz = __fdiv_rn(bz, s); }
    else { qx = bx * inv_s; qy = by * inv_s; qz = bz * inv_s; }
}

// ---- binary operators: a = left operand (deeper stack entry), b = right (top) ----------------
// X2: h = max(k-|a-b|,0)/k ; min(a,b) - h*h*k*0.25
// Exact mode evaluates that literally (inv_k = 1/k, used when k is a power of two).  Fast mode uses the algebraically
// equal  min(a,b) - e*e,  e = max(k*s - |a-b|*s, 0),  s = sqrt(1/(4k)):  the FAST form of the program carries
// (k*s, s) in the (inv_k, c4k) positions (rm_program.h smooth_constants_fast), and the blend is FADD, FFMA, 2 FMNMX, FFMA.
// SAT (fast mode, k*s <= 1 — rm_program.h smooth_fast_sat_ok): the clamp at 0 as the FFMA's .SAT modifier, same value.
template <bool EX, bool SAT = false>
__device__ __forceinline__ float op_smooth(float a, float b, float k, float inv_k, float c4k, bool k_pow2) {
    using A = Arith<EX>;
    if (EX) {
        float e = fmaxf(A::sub(k, fabsf(A::sub(a, b))), 0.0f);
        float m = fminf(a, b);
        float h = k_pow2 ? __fmul_rn(e, inv_k) : __fdiv_rn(e, k);
        return __fsub_rn(m, __fmul_rn(__fmul_rn(__fmul_rn(h, h), k), 0.25f));
    } else {
        const float ks = inv_k, s = c4k;
        const float e = SAT ? __saturatef(fmaf(-fabsf(a - b), s, ks)) : fmaxf(fmaf(-fabsf(a - b), s, ks), 0.0f);
        return fmaf(-e, e, fminf(a, b));
    }
}
// acc = smooth-min(acc, b), same operations, with the result tied to acc's register so that the
// interpreter loop carries no accumulator copies around its back-edge.
template <bool EX>
__device__ __forceinline__ void op_smooth_acc(float &acc, float b, float k, float inv_k, float c4k, bool k_pow2) {
    using A = Arith<EX>;
    if (EX) {
        const float e = fmaxf(A::sub(k, fabsf(A::sub(acc, b))), 0.0f);
        RM_PTX_MIN_ACC(acc, b);
        const float h = k_pow2 ? __fmul_rn(e, inv_k) : __fdiv_rn(e, k);
        const float t = __fmul_rn(__fmul_rn(__fmul_rn(h, h), k), 0.25f);
        RM_PTX_SUB_ACC(acc, t);
    } else {
        const float ks = inv_k, s = c4k;
        const float e = fmaxf(fmaf(-fabsf(acc - b), s, ks), 0.0f);
        RM_PTX_MIN_ACC(acc, b);
        const float t = -e;
        RM_PTX_FMA_ACC(acc, t, e);
    }
}
template <bool EX>
__device__ __forceinline__ float combine(uint32_t cb, float a, float b, const float4 &q1, float c4k, bool k_pow2) {
    switch (cb) {
        case CB_UNION: return fminf(a, b);   // wgsl:242-246
        case CB_SUB: return fmaxf(a, -b);    // wgsl:248-252
        case CB_INTER: return fmaxf(a, b);   // X1
        default: return op_smooth<EX>(a, b, q1.z, q1.w, c4k, k_pow2);
    }
}

// ---- map_scene for one point: the decoded-program interpreter (wgsl:187-203) -----------------
// prog: decoded instructions (shared or global memory), viewed as float4[3*n].
// EXT: the program may hold N4 instructions (rm_program.h): the primitives then read the evaluation point q that the
// XOP_* instructions maintain.  EXT = false compiles to the reference-set interpreter, unchanged.
template <bool EX, bool EXT = false>
__device__ __forceinline__ float map_eval1(const float4 *__restrict__ prog, uint32_t n_insn, float max_dist,
                                           float px, float py, float pz) {
    if (n_insn == 0) return max_dist;  // wgsl:189-191
    float acc = 0.0f;
    float spill[RM_STACK_MAX - 1];
    float qx = px, qy = py, qz = pz;
    for (uint32_t i = 0; i < n_insn; ++i) {
        const float4 q0 = prog[3 * i], q1 = prog[3 * i + 1], q2 = prog[3 * i + 2];
        const uint32_t op = __float_as_uint(q2.y), slot = __float_as_uint(q2.z);
        if (EXT) {
            const uint32_t fl = __float_as_uint(q2.w);
            if (fl & FLAG_ACCMUL) {  // ScalePop
                acc = Arith<EX>::mul(acc, q0.x);
                continue;
            }
            if (fl & FLAG_XF) {
                const bool world = (slot & 1u) != 0;  // first step of a chain starts from the ray position
                const float bx = world ? px : qx, by = world ? py : qy, bz = world ? pz : qz;
                if (op == XOP_TRANSLATE) xf_translate<EX>(bx, by, bz, q0.x, q0.y, q0.z, qx, qy, qz);
                else if (op == XOP_ROTATE) xf_rotate<EX>(bx, by, bz, q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w, q2.x, qx, qy, qz);
                else if (op == XOP_SCALE) xf_scale<EX>(bx, by, bz, q0.x, q0.y, qx, qy, qz);
                else { qx = px; qy = py; qz = pz; }  // XOP_RESET
                continue;
            }
        }
        const float ex = EXT ? qx : px, ey = EXT ? qy : py, ez = EXT ? qz : pz;
        const bool k_pow2 = (__float_as_uint(q2.w) & FLAG_K_POW2) != 0;
        const uint32_t src = EXT ? opcode_src_ext(op) : opcode_src(op), cb = EXT ? opcode_cb_ext(op) : opcode_cb(op);
        float a, b;
        if (src == SRC_SPHERE) {
            b = sd_sphere<EX>(ex, ey, ez, q0.x, q0.y, q0.z, q0.w);
            a = acc;
        } else if (src == SRC_BOX) {
            b = sd_box<EX>(ex, ey, ez, q0.x, q0.y, q0.z, q0.w, q1.x, q1.y);
            a = acc;
        } else if (EXT && src == SRC_PLANE) {
            b = sd_plane<EX>(ex, ey, ez, q0.x, q0.y, q0.z, q0.w);
            a = acc;
        } else {
            b = acc;
            a = spill[slot];
        }
        if (cb == CB_FIRST) acc = b;
        else if (cb == CB_PUSH) { spill[slot] = acc; acc = b; }
        else acc = combine<EX>(cb, a, b, q1, q2.x, k_pow2);
    }
    return acc;
}

// ---- ray generation: fs_main wgsl:39-62 with the pixel->pt_screen closed form --
__device__ __forceinline__ void mat4_mul(const float *m, float x, float y, float z, float w, float out[4]) {
    // column sum accumulated left to right, one rounding per op (SURVEY.md §8(c).1 rule 3)
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        float acc = __fmul_rn(m[r], x);
        acc = __fadd_rn(acc, __fmul_rn(m[4 + r], y));
        acc = __fadd_rn(acc, __fmul_rn(m[8 + r], z));
        acc = __fadd_rn(acc, __fmul_rn(m[12 + r], w));
        out[r] = acc;
    }
}
// Matrix * vector in the arithmetic mode of the caller: exact = one rounding per operation, left to right (rule 3).
template <bool EX>
__device__ __forceinline__ void mat4_mul_t(const float *m, float x, float y, float z, float w, float out[4]) {
    if (EX) {
        mat4_mul(m, x, y, z, w, out);
    } else {
#pragma unroll
        for (int r = 0; r < 4; ++r) out[r] = fmaf(m[12 + r], w, fmaf(m[8 + r], z, fmaf(m[4 + r], y, m[r] * x)));
    }
}
// Ray generation (wgsl:39-62).  Exact mode evaluates it with one rounding per operation.  Fast mode uses FMA matrix
// products and a reciprocal square root: for a 3-instruction program ray generation is ~20 % of the frame
// (profiles/r1_s6_NOTES.md), not the ~0.5 % it is for the deep scene.  The pixel -> pt_screen mapping stays exact in
// both modes (one division per pixel), and so do the per-frame constants in FrameParams.
template <bool EX>
__device__ __forceinline__ void gen_ray(const FrameParams &fp, uint32_t px, uint32_t py, uint32_t sample,
                                        float ro[3], float rd[3]) {
    const float sx = __fsub_rn(__fmul_rn(__fdiv_rn(__fadd_rn((float)px, 0.5f), (float)fp.W), 2.0f), 1.0f);
    const float sy = __fsub_rn(1.0f, __fmul_rn(__fdiv_rn(__fadd_rn((float)py, 0.5f), (float)fp.H), 2.0f));
    const float ax = fp.sample_ax[sample >> 2], ay = fp.sample_ay[sample & 3u];  // wgsl:44-45: i outer (x), j inner (y)
    float pv[4], pw[4];
    mat4_mul_t<EX>(fp.inv_proj, __fadd_rn(sx, ax), __fadd_rn(sy, ay), -1.0f, 1.0f, pv);  // wgsl:56-58
    mat4_mul_t<EX>(fp.inv_view, pv[0], pv[1], pv[2], pv[3], pw);                          // wgsl:59
    const float dx = __fsub_rn(pw[0], fp.ro4[0]), dy = __fsub_rn(pw[1], fp.ro4[1]);
    const float dz = __fsub_rn(pw[2], fp.ro4[2]), dw = __fsub_rn(pw[3], fp.ro4[3]);
    if (EX) {
        const float len = __fsqrt_rn(__fadd_rn(
            __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)), __fmul_rn(dw, dw)));
        rd[0] = __fdiv_rn(dx, len); rd[1] = __fdiv_rn(dy, len); rd[2] = __fdiv_rn(dz, len);  // wgsl:62
    } else {
        const float inv = rsqrtf(fmaf(dw, dw, fmaf(dz, dz, fmaf(dy, dy, dx * dx))));
        rd[0] = dx * inv; rd[1] = dy * inv; rd[2] = dz * inv;
    }
    ro[0] = fp.ro4[0]; ro[1] = fp.ro4[1]; ro[2] = fp.ro4[2];
}

// Fast-mode ray generation from the per-frame affine form (FrameParams::ray_c0 / ray_cx / ray_cy): same rays as gen_ray<false>
// up to float rounding (the composition is done in double on the host), a third of the instructions.
__device__ __forceinline__ void gen_ray_affine(const FrameParams &fp, float fpx, float fpy, uint32_t sample, float rd[3]) {
    const float X = fmaf((float)(sample >> 2), fp.ray_gx, fpx), Y = fmaf((float)(sample & 3u), fp.ray_gy, fpy);  // wgsl:44-45: i -> x, j -> y
    const float dx = fmaf(Y, fp.ray_cy[0], fmaf(X, fp.ray_cx[0], fp.ray_c0[0]));
    const float dy = fmaf(Y, fp.ray_cy[1], fmaf(X, fp.ray_cx[1], fp.ray_c0[1]));
    const float dz = fmaf(Y, fp.ray_cy[2], fmaf(X, fp.ray_cx[2], fp.ray_c0[2]));
    const float dw = fmaf(Y, fp.ray_cy[3], fmaf(X, fp.ray_cx[3], fp.ray_c0[3]));
    const float inv = rsqrtf(fmaf(dw, dw, fmaf(dz, dz, fmaf(dy, dy, dx * dx))));  // vec4 normalise, wgsl:62
    rd[0] = dx * inv; rd[1] = dy * inv; rd[2] = dz * inv;
}

// ---- hit shading wgsl:100-105 given the un-normalised tetrahedron sum n (wgsl:138-143) --------
template <bool EX>
__device__ __forceinline__ float shade_diffuse(float nx, float ny, float nz, float px, float py, float pz) {
    using A = Arith<EX>;
    A::normalize3(nx, ny, nz);
    float lx = A::sub(px, RM_LIGHT_X), ly = A::sub(py, RM_LIGHT_Y), lz = A::sub(pz, RM_LIGHT_Z);
    A::normalize3(lx, ly, lz);
    float d = A::add(A::add(A::mul(nx, lx), A::mul(ny, ly)), A::mul(nz, lz));
    return fmaxf(RM_AMBIENT, d);
}

// normal from the four tetrahedron taps + diffuse term (wgsl:98-105, 138-143) -> sample code.  A rare per-ray event,
// kept OUT OF LINE (one copy) so that the marching loops keep a small instruction footprint.
template <bool EX>
__device__ __noinline__ float wf_shade(float a0, float a1, float a2, float a3, float hx, float hy, float hz) {
    using A = Arith<EX>;
    const float nx = A::add(A::add(A::add(a0, -a1), -a2), a3);
    const float ny = A::add(A::add(A::add(-a0, -a1), a2), a3);
    const float nz = A::add(A::add(A::add(-a0, a1), -a2), a3);
    return shade_diffuse<EX>(nx, ny, nz, hx, hy, hz);
}

// ---- floor / sky wgsl:117-130.  Returns class; colour via c[3]; t = floor_dist. ---------------
__device__ __forceinline__ int shade_miss(const float ro[3], const float rd[3], float c[3], float &t) {
    const float fd = __fdiv_rn(__fsub_rn(RM_FLOOR_Y, ro[1]), rd[1]);  // wgsl:120
    t = fd;
    if (fd > 0.0f) {
        const float fx = __fadd_rn(ro[0], __fmul_rn(rd[0], fd)), fz = __fadd_rn(ro[2], __fmul_rn(rd[2], fd));
        // round = RNE (rintf); f32->i32 saturating, NaN -> 0: cvt.rni.s32.f32 does both (Q6/Q7)
        const int ix = __float2int_rn(__fadd_rn(fx, 0.5f)), iz = __float2int_rn(__fadd_rn(fz, 0.5f));
        const float col = (float)((ix ^ iz) & 1);
        c[0] = __fadd_rn(0.1f, __fmul_rn(0.2f, col));
        c[1] = c[0];
        c[2] = __fadd_rn(0.2f, __fmul_rn(0.2f, col));
        return 1;
    }
    c[0] = c[1] = c[2] = 0.0f;
    return 0;
}

// unorm8 = floor(clamp(v,0,1)*255 + 0.5)
__device__ __forceinline__ uint32_t unorm8(float v) {
    float c = fminf(fmaxf(v, 0.0f), 1.0f);  // NaN -> 0 via fmaxf
    return (uint32_t)floorf(__fadd_rn(__fmul_rn(c, 255.0f), 0.5f));
}

// Per-sample result code written by the marching kernels and consumed by the resolve:
//   v >= 0   object hit, v = diffuse intensity (>= 0.02)      wgsl:103-105
//   v == -1  floor, dark cell;  v == -2  floor, light cell   wgsl:124-127
//   v == -3  sky                                              wgsl:130
#define RM_CODE_FLOOR_DARK (-1.0f)
#define RM_CODE_FLOOR_LIGHT (-2.0f)
#define RM_CODE_SKY (-3.0f)

// shade_miss, returning the sample code directly (the cell's parity picks the code; no colour is formed).
__device__ __forceinline__ float shade_miss_code(const float ro[3], const float rd[3], float &t, int &cls) {
    const float fd = __fdiv_rn(__fsub_rn(RM_FLOOR_Y, ro[1]), rd[1]);  // wgsl:120
    t = fd;
    cls = 0;
    if (fd > 0.0f) {
        const float fx = __fadd_rn(ro[0], __fmul_rn(rd[0], fd)), fz = __fadd_rn(ro[2], __fmul_rn(rd[2], fd));
        const int ix = __float2int_rn(__fadd_rn(fx, 0.5f)), iz = __float2int_rn(__fadd_rn(fz, 0.5f));
        cls = 1;
        return ((ix ^ iz) & 1) ? RM_CODE_FLOOR_LIGHT : RM_CODE_FLOOR_DARK;  // 0.1 + 0.2 col > 0.2 iff col == 1
    }
    return RM_CODE_SKY;
}

#define RM_GAMMA_FLOOR_DARK_RG 0.31622776f   /* 0x3ea1e89b */
#define RM_GAMMA_FLOOR_DARK_B 0.4472136f     /* 0x3ee4f92e */
#define RM_GAMMA_FLOOR_LIGHT_RG 0.5477226f   /* 0x3f0c378c */
#define RM_GAMMA_FLOOR_LIGHT_B 0.6324555f    /* 0x3f21e89b */

// Colour of a sample code, gamma'd (wgsl:68) — always with exact roundings (negligible cost).
__device__ __forceinline__ void code_to_gamma(float code, float g[3]) {
    if (code >= 0.0f) {
        g[0] = __fsqrt_rn(__fmul_rn(RM_ALBEDO_R, code));
        g[1] = __fsqrt_rn(__fmul_rn(RM_ALBEDO_G, code));
        g[2] = __fsqrt_rn(__fmul_rn(RM_ALBEDO_B, code));
    } else if (code == RM_CODE_SKY) {
        g[0] = g[1] = g[2] = 0.0f;
    } else {
        // sqrt_rn(0.1 + 0.2 col), sqrt_rn(0.2 + 0.2 col) (wgsl:124-127, 68) for col = 0 / 1: four constants, the correctly
        // rounded roots of fl(0.1), fl(0.2), fl(0.1 + 0.2), fl(0.2 + 0.2) (tests/test_hostcheck_cpu.py recomputes them)
        const bool light = code == RM_CODE_FLOOR_LIGHT;
        g[0] = g[1] = light ? RM_GAMMA_FLOOR_LIGHT_RG : RM_GAMMA_FLOOR_DARK_RG;
        g[2] = light ? RM_GAMMA_FLOOR_LIGHT_B : RM_GAMMA_FLOOR_DARK_B;
    }
}

// The same in fast mode: sqrt(albedo * v) as sqrt(albedo) * sqrt(v) — one MUFU and three multiplies instead of three IEEE
// square roots (a relative difference of ~1e-7 per sample before the 8-bit quantisation).  Floor and sky samples are the
// exact constants in both modes.
#define RM_SQRT_ALBEDO_R 0.632455532f   /* sqrt(0.4) */
#define RM_SQRT_ALBEDO_G 0.836660027f   /* sqrt(0.7) */
#define RM_SQRT_ALBEDO_B 0.316227766f   /* sqrt(0.1) */
template <bool EX>
__device__ __forceinline__ void code_to_gamma_t(float code, float g[3]) {
    if (EX) {
        code_to_gamma(code, g);
    } else if (code >= 0.0f) {
        float r;
        RM_PTX_SQRT_APPROX(r, code);
        g[0] = RM_SQRT_ALBEDO_R * r; g[1] = RM_SQRT_ALBEDO_G * r; g[2] = RM_SQRT_ALBEDO_B * r;
    } else {
        code_to_gamma(code, g);
    }
}

// sum / 16 (wgsl:75) as sum * 0.0625: 16 is a power of two, so both are the correctly rounded value of the same real
// number for every float (subnormal results, infinities and NaN included) — one multiply instead of an IEEE division.
__device__ __forceinline__ uint32_t pack_mean16(float sr, float sg, float sb) {
    return unorm8(__fmul_rn(sr, 0.0625f)) | (unorm8(__fmul_rn(sg, 0.0625f)) << 8) | (unorm8(__fmul_rn(sb, 0.0625f)) << 16) |
           0xff000000u;
}

// In-order 16-sample mean (wgsl:44-45,68-75) and RGBA8 pack.
__device__ __forceinline__ uint32_t resolve_pixel(const float *codes16) {
    float sr = 0.0f, sg = 0.0f, sb = 0.0f;
#pragma unroll
    for (int s = 0; s < 16; ++s) {
        float g[3];
        code_to_gamma(codes16[s], g);
        sr = __fadd_rn(sr, g[0]); sg = __fadd_rn(sg, g[1]); sb = __fadd_rn(sb, g[2]);
    }
    return pack_mean16(sr, sg, sb);
}

}  // namespace rm
