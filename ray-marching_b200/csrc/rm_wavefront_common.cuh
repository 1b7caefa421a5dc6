// rm_wavefront_common.cuh — pieces shared by the two persistent wavefront kernels
// (rm_wavefront.cu: per-slot state machine; rm_wavefront2.cu: march slots + hit queue):
// tile geometry, ray-slot meta word, the R-wide decoded-program interpreter (wgsl:187-252),
// and the out-of-line shade / resolve helpers (wgsl:98-105,138-143 and wgsl:44-45,68-75).
#pragma once
#include "rm_wavefront.h"

namespace rm {
namespace {

constexpr int WF_WARPS = 4;      // warps per CTA
constexpr int WF_K = 4;          // tiles in flight per warp
constexpr int TILE_W = 8, TILE_H = 4;
constexpr int TILE_RAYS = TILE_W * TILE_H * 16;  // 512
constexpr unsigned FULL = 0xffffffffu;
constexpr int RING = 1024;

constexpr uint32_t M_Q = 0x1ffu;          // sample slot in tile: sample*32 + pixel
constexpr uint32_t M_BUF_SHIFT = 9;       // 2 bits
constexpr uint32_t M_STATE_SHIFT = 11;    // 3 bits: 0 MARCH, 1..4 TAP0..TAP3
constexpr uint32_t M_LIVE = 1u << 14;
constexpr uint32_t M_ITER_SHIFT = 16;      // 16 bits: march iterations done (max_iter <= 65535)

constexpr int RING_N = 64;  // ready-ray ring: two batches of 32
// map_scene for R points at once: one instruction fetch and ONE jump-table dispatch on the fused
// opcode feed R evaluations; every case folds its source straight into the accumulators.
template <bool EX, uint32_t SRC>
__device__ __forceinline__ float prim_eval(float x, float y, float z, const float4 &q0, const float4 &q1) {
    if (SRC == SRC_SPHERE) return sd_sphere<EX>(x, y, z, q0.x, q0.y, q0.z, q0.w);
    return sd_box<EX>(x, y, z, q0.x, q0.y, q0.z, q0.w, q1.x, q1.y);
}

#define RM_OPC(SRC, CB) opcode(SRC, CB)
#define RM_PRIM_CASES(SRC)                                                                                  \
    case RM_OPC(SRC, CB_FIRST):                                                                             \
        RM_FOR_R acc[r] = prim_eval<EX, SRC>(px[r], py[r], pz[r], q0, q1);                                  \
        break;                                                                                              \
    case RM_OPC(SRC, CB_PUSH):                                                                              \
        RM_FOR_R { spill[slot][r] = acc[r]; acc[r] = prim_eval<EX, SRC>(px[r], py[r], pz[r], q0, q1); }     \
        break;                                                                                              \
    case RM_OPC(SRC, CB_UNION):                                                                             \
        RM_FOR_R acc[r] = fminf(acc[r], prim_eval<EX, SRC>(px[r], py[r], pz[r], q0, q1));                   \
        break;                                                                                              \
    case RM_OPC(SRC, CB_SUB):                                                                               \
        RM_FOR_R acc[r] = fmaxf(acc[r], -prim_eval<EX, SRC>(px[r], py[r], pz[r], q0, q1));                  \
        break;                                                                                              \
    case RM_OPC(SRC, CB_INTER):                                                                             \
        RM_FOR_R acc[r] = fmaxf(acc[r], prim_eval<EX, SRC>(px[r], py[r], pz[r], q0, q1));                   \
        break;                                                                                              \
    case RM_OPC(SRC, CB_SMOOTH):                                                                            \
        RM_FOR_R acc[r] = op_smooth<EX>(acc[r], prim_eval<EX, SRC>(px[r], py[r], pz[r], q0, q1), q1.z, q1.w, \
                                        q2.x, k_pow2);                                                      \
        break;

// acc = min(acc, v) etc. with the accumulator tied to one register (no copies around the loop)
__device__ __forceinline__ void acc_min(float &acc, float v) { RM_PTX_MIN_ACC(acc, v); }
__device__ __forceinline__ void acc_max(float &acc, float v) { RM_PTX_MAX_ACC(acc, v); }

// Early-out (EO, opt-in, off in every default path): for acc = op(acc, prim(p)) with op in {union, smooth-union,
// subtraction} the accumulator is provably unchanged — bit for bit, in both arithmetic modes — when a cheap lower bound
// of prim(p) already exceeds  acc (+ k for the smooth union), resp. -acc for the subtraction:
//   sphere: |p-c|^2 >= T*max(T,0) with T = +-acc + r (+ k) + slack;   box: max(qx,qy,qz) >= +-acc (+ k) + slack
// (max(q) <= box distance).  Then e = max(k - |acc - v|, 0) is exactly 0 and min(acc, v) = acc / max(acc, -v) = acc.
// If that holds for all R rays of the thread the rest of the instruction is skipped (sqrt, the blend, most of the box).
// The slack (1e-4) covers the few roundings between the bound and the value for coordinates up to max_dist = 100.
#define RM_EO_SLACK 1e-4f
// EXT (N4, rm_program.h): the program may hold Plane primitives and the XOP_* instructions that maintain the
// evaluation point q of the R rays; primitives then read q.  EXT = false is the reference-set interpreter, unchanged.
// ---- the fast-mode interpreter of the reference instruction set (map_evalR_tail) --------------------------------------
// COMMON TAIL: every dispatch path leaves (m, e) per ray and ONE tail writes acc = m - e*e (e = 0 for union / subtraction /
// intersection / first / push: the value is unchanged bit for bit).  The accumulators then have a single definition per
// trip and ptxas keeps them in place; in map_evalR below the results of the taken path are copied back to the loop's
// accumulator registers at the back-edge (8 MOV per instruction with 8 rays).  The smooth union needs the FFMA anyway.
// Measured on B200 (profiles/r2_interp_NOTES.md): cfg3 with 8 rays 42.8 -> 41.2 ms; with 4 rays (short programs) it
// loses (cfg2 1.07 -> 1.20 ms), so only the 8-ray kernels use it.  Dispatching PAIRS of instructions at once (decoder-
// marked, 8 templated pair bodies) and 16 rays per thread were measured too and lost: 44.4 and 42.0 ms.
// Evaluating the 8 rays as 4 PAIRS through FADD2 / FMUL2 / FFMA2 (rm_device.cuh f2_t — what the compiled kernels do) was
// measured here too: 38.74 against 38.55 ms, no gain — at 72 registers the 64-bit register pairs cost ~350 MOV / IMAD.MOV
// in the kernel and 192 instead of 144 bytes of spill (profiles/r2_s45_packed.txt).
template <int R, bool BOX>
__device__ __forceinline__ void tail_prim(const float4 &q0, const float4 &q1, const float (&px)[R], const float (&py)[R],
                                          const float (&pz)[R], float (&v)[R]) {
    if (BOX) { RM_FOR_R v[r] = sd_box<false>(px[r], py[r], pz[r], q0.x, q0.y, q0.z, q0.w, q1.x, q1.y); }
    else { RM_FOR_R v[r] = sd_sphere<false>(px[r], py[r], pz[r], q0.x, q0.y, q0.z, q0.w); }
}

template <int R>
__device__ __forceinline__ void map_evalR_tail(const float4 *__restrict__ prog, uint32_t n_insn, float max_dist,
                                               const float (&px)[R], const float (&py)[R], const float (&pz)[R],
                                               float (&acc)[R], float (*spill)[R]) {
    if (n_insn == 0) {  // wgsl:189-191
        RM_FOR_R acc[r] = max_dist;
        return;
    }
    RM_FOR_R acc[r] = 0.0f;
#pragma unroll 1
    for (uint32_t i = 0; i < n_insn; ++i) {
        const float4 q0 = prog[3 * i], q1 = prog[3 * i + 1], q2 = prog[3 * i + 2];
        const uint32_t fl = __float_as_uint(q2.w);
        float m[R], e[R];
        if (!(fl & FLAG_POP)) {
            float v[R];
            if (fl & FLAG_BOX) tail_prim<R, true>(q0, q1, px, py, pz, v);
            else tail_prim<R, false>(q0, q1, px, py, pz, v);
            if (fl & FLAG_SMOOTH) {
                if (fl & FLAG_SAT_OK) {  // k*s <= 1: the clamp at 0 is the FFMA's .SAT (same value, one FMNMX fewer per ray)
                    RM_FOR_R { e[r] = __saturatef(fmaf(-fabsf(acc[r] - v[r]), q2.x, q1.w)); m[r] = fminf(acc[r], v[r]); }
                } else {
                    RM_FOR_R { e[r] = fmaxf(fmaf(-fabsf(acc[r] - v[r]), q2.x, q1.w), 0.0f); m[r] = fminf(acc[r], v[r]); }
                }
            } else {
                if (fl & FLAG_UNION) {
                    RM_FOR_R m[r] = fminf(acc[r], v[r]);
                } else if (fl & FLAG_SUB) {
                    RM_FOR_R m[r] = fmaxf(acc[r], -v[r]);
                } else if (fl & FLAG_INTER) {
                    RM_FOR_R m[r] = fmaxf(acc[r], v[r]);
                } else {
                    if (fl & FLAG_PUSH) {
                        const uint32_t slot = __float_as_uint(q2.z);
                        RM_FOR_R spill[slot][r] = acc[r];
                    }
                    RM_FOR_R m[r] = v[r];
                }
                RM_FOR_R e[r] = 0.0f;
            }
        } else {
            const uint32_t slot = __float_as_uint(q2.z);
            if (fl & FLAG_SMOOTH) {
                RM_FOR_R {
                    const float a = spill[slot][r];
                    e[r] = fmaxf(fmaf(-fabsf(a - acc[r]), q2.x, q1.w), 0.0f);
                    m[r] = fminf(a, acc[r]);
                }
            } else {
                if (fl & FLAG_UNION) {
                    RM_FOR_R m[r] = fminf(spill[slot][r], acc[r]);
                } else if (fl & FLAG_SUB) {
                    RM_FOR_R m[r] = fmaxf(spill[slot][r], -acc[r]);
                } else {
                    RM_FOR_R m[r] = fmaxf(spill[slot][r], acc[r]);
                }
                RM_FOR_R e[r] = 0.0f;
            }
        }
        RM_FOR_R acc[r] = fmaf(-e[r], e[r], m[r]);
    }
}

template <bool EX, int R, bool EO = false, bool EXT = false>
__device__ __forceinline__ void map_evalR(const float4 *__restrict__ prog, uint32_t n_insn, float max_dist,
                                          const float (&px)[R], const float (&py)[R], const float (&pz)[R],
                                          float (&acc)[R], float (*spill)[R]) {
    static_assert(!(EO && EXT), "the early-out bounds are written for the reference instruction set");
    if (n_insn == 0) {  // wgsl:189-191
        RM_FOR_R acc[r] = max_dist;
        return;
    }
    RM_FOR_R acc[r] = 0.0f;
    float qx[R], qy[R], qz[R];  // EXT only: evaluation point in force (starts as the ray position)
    if (EXT) { RM_FOR_R { qx[r] = px[r]; qy[r] = py[r]; qz[r] = pz[r]; } }
#define RM_EVX (EXT ? qx[r] : px[r])
#define RM_EVY (EXT ? qy[r] : py[r])
#define RM_EVZ (EXT ? qz[r] : pz[r])
    // not unrolled: unrolling by 2 removes the 8 accumulator copies ptxas leaves at the back-edge, but measured
    // slower on B200 (cfg3 48.94 vs 48.64 ms, cfg2 2.77 vs 2.32 ms: larger code, remainder loop) — profiles/r1_s2_NOTES.md
#pragma unroll 1
    for (uint32_t i = 0; i < n_insn; ++i) {
        const float4 q0 = prog[3 * i], q1 = prog[3 * i + 1], q2 = prog[3 * i + 2];
        const uint32_t fl = __float_as_uint(q2.w);
        if (EXT && (fl & (FLAG_XF | FLAG_ACCMUL))) {
            if (fl & FLAG_ACCMUL) {  // ScalePop: the child's distance back in the parent's units
                RM_FOR_R acc[r] = Arith<EX>::mul(acc[r], q0.x);
                continue;
            }
            const uint32_t xop = __float_as_uint(q2.y);
            if (__float_as_uint(q2.z) & 1u) { RM_FOR_R { qx[r] = px[r]; qy[r] = py[r]; qz[r] = pz[r]; } }  // chain starts at p
            if (xop == XOP_TRANSLATE) {
                RM_FOR_R xf_translate<EX>(qx[r], qy[r], qz[r], q0.x, q0.y, q0.z, qx[r], qy[r], qz[r]);
            } else if (xop == XOP_ROTATE) {
                RM_FOR_R xf_rotate<EX>(qx[r], qy[r], qz[r], q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w, q2.x, qx[r], qy[r], qz[r]);
            } else if (xop == XOP_SCALE) {
                RM_FOR_R xf_scale<EX>(qx[r], qy[r], qz[r], q0.x, q0.y, qx[r], qy[r], qz[r]);
            }  // XOP_RESET: the copy above is all of it
            continue;
        }
        // Dispatch on one-hot flag bits, most frequent fused ops first (primitive kind, then
        // smooth-union / union / subtraction / intersection / push / first).
        if (!(fl & FLAG_POP)) {
            if (EO && (fl & (FLAG_SMOOTH | FLAG_UNION | FLAG_SUB))) {
                using A = Arith<EX>;
                const float sgn = (fl & FLAG_SUB) ? -1.0f : 1.0f;
                const float thr = ((fl & FLAG_SMOOTH) ? q1.z : 0.0f) + RM_EO_SLACK;
                // one signed margin per ray, all independent (ILP = R), then ONE minimum and ONE branch: written with
                // `skip = skip && cond` the compiler short-circuits ray by ray into a serial chain of predicated tests
                float worst = 3.0e38f;
                if (fl & FLAG_BOX) {
                    RM_FOR_R {  // margin = max(q) - T,  T = +-acc (+ k) + slack;  max(q) <= box distance
                        const float qx = A::sub(fabsf(A::sub(px[r], q0.x)), q0.w);
                        const float qy = A::sub(fabsf(A::sub(py[r], q0.y)), q1.x);
                        const float qz = A::sub(fabsf(A::sub(pz[r], q0.z)), q1.y);
                        worst = fminf(worst, fmaxf(qx, fmaxf(qy, qz)) - fmaf(sgn, acc[r], thr));
                    }
                } else {
                    const float thr_s = thr + q0.w;
                    RM_FOR_R {  // margin = |p-c|^2 - T*max(T,0),  T = +-acc + r (+ k) + slack.  T < 0 means the accumulator is
                                // already below -r <= sphere distance, so the margin |p-c|^2 >= 0 is right there too
                        const float dx = A::sub(px[r], q0.x), dy = A::sub(py[r], q0.y), dz = A::sub(pz[r], q0.z);
                        const float t = fmaf(sgn, acc[r], thr_s);
                        worst = fminf(worst, fmaf(-t, fmaxf(t, 0.0f), A::dot3s(dx, dy, dz)));
                    }
                }
                if (worst >= 0.0f) continue;
            }
            float v[R];
            if (fl & FLAG_BOX) {
                RM_FOR_R v[r] = sd_box<EX>(RM_EVX, RM_EVY, RM_EVZ, q0.x, q0.y, q0.z, q0.w, q1.x, q1.y);
            } else if (EXT && (fl & FLAG_PLANE)) {
                RM_FOR_R v[r] = sd_plane<EX>(RM_EVX, RM_EVY, RM_EVZ, q0.x, q0.y, q0.z, q0.w);
            } else {
                RM_FOR_R v[r] = sd_sphere<EX>(RM_EVX, RM_EVY, RM_EVZ, q0.x, q0.y, q0.z, q0.w);
            }
            if (fl & FLAG_SMOOTH) {
                RM_FOR_R op_smooth_acc<EX>(acc[r], v[r], q1.z, q1.w, q2.x, (fl & FLAG_K_POW2) != 0);
            } else if (fl & FLAG_UNION) {
                RM_FOR_R acc_min(acc[r], v[r]);
            } else if (fl & FLAG_SUB) {
                RM_FOR_R acc_max(acc[r], -v[r]);
            } else if (fl & FLAG_INTER) {
                RM_FOR_R acc_max(acc[r], v[r]);
            } else {
                if (fl & FLAG_PUSH) {
                    const uint32_t slot = __float_as_uint(q2.z);
                    RM_FOR_R spill[slot][r] = acc[r];
                }
                RM_FOR_R acc[r] = v[r];
            }
        } else {
            // binary operator on two subtrees: right operand = accumulator, left = spilled entry
            const uint32_t slot = __float_as_uint(q2.z);
            if (fl & FLAG_SMOOTH) {
                RM_FOR_R acc[r] = op_smooth<EX>(spill[slot][r], acc[r], q1.z, q1.w, q2.x, (fl & FLAG_K_POW2) != 0);
            } else if (fl & FLAG_UNION) {
                RM_FOR_R acc[r] = fminf(spill[slot][r], acc[r]);
            } else if (fl & FLAG_SUB) {
                RM_FOR_R acc[r] = fmaxf(spill[slot][r], -acc[r]);
            } else {
                RM_FOR_R acc[r] = fmaxf(spill[slot][r], acc[r]);
            }
        }
    }
#undef RM_EVX
#undef RM_EVY
#undef RM_EVZ
}

// map_scene through the decoded-program interpreter (program in shared memory, spill slots in the thread's frame)
// The 8-ray fast-mode kernels of reference-set programs run map_evalR_tail; exact mode, the early-out, extended programs
// and the 4-ray kernels run map_evalR.
template <bool EX, int R, bool EO, bool EXT>
struct InterpMap {
    const float4 *prog;
    uint32_t n_insn;
    float max_dist;
    float spill[RM_STACK_MAX - 1][R];
    __device__ __forceinline__ void operator()(const float (&px)[R], const float (&py)[R], const float (&pz)[R], float (&s)[R]) {
        if constexpr (R == 8 && !EX && !EO && !EXT) map_evalR_tail<R>(prog, n_insn, max_dist, px, py, pz, s, spill);
        else map_evalR<EX, R, EO, EXT>(prog, n_insn, max_dist, px, py, pz, s, spill);
    }
};

// ---- rare per-ray events, kept OUT OF LINE (one copy each) so that the per-iteration instruction
// footprint of the persistent loop stays inside the instruction cache (profiles/r1_v1_*: the fully
// inlined R=4 kernel stalled 4.7 cycles per issue on no_instruction) ------------------------------
// (wf_shade, the normal + diffuse term of a hit, lives in rm_device.cuh: the direct kernel shares it)

// in-order 16-sample resolve of one pixel of a finished tile (codes[sample*32 + pixel])
__device__ __noinline__ uint32_t wf_resolve(const float *codes, int lane) {
    float sr = 0.0f, sg = 0.0f, sb = 0.0f;
#pragma unroll 1
    for (int sidx = 0; sidx < 16; ++sidx) {
        float g[3];
        code_to_gamma(codes[sidx * 32 + lane], g);
        sr = __fadd_rn(sr, g[0]); sg = __fadd_rn(sg, g[1]); sb = __fadd_rn(sb, g[2]);
    }
    return pack_mean16(sr, sg, sb);
}

}  // namespace
}  // namespace rm
