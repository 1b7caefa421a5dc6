// rm_program.h — host-side validation and pre-decoding of the reference's postfix CSG command
// stream (wire format: /root/reference/src/ray_marching/csg/builder.rs:26-62; interpreter it
// replaces: ray_marching.wgsl:146-252) into the fixed-width instruction stream the kernels run.
//
// The program is identical for every ray, so everything that the WGSL interpreter resolves at
// run time per thread — cursor arithmetic, opcode decode, value-stack indices — is resolved here
// once per rm_prepare:
//   * the top of the value stack lives in a register (the accumulator); deeper entries live in
//     statically numbered spill slots;
//   * "primitive immediately followed by a binary operator" is fused into one instruction
//     acc = op(acc, prim(p)), which is the whole program for left-deep chains;
//   * per-instruction constants that are exact (k*0.25, 1/k when k is a power of two) are
//     precomputed.  Nothing here changes evaluation order or rounding.
//
// N4 extension (include/rm_b200.h: Plane, Translation / Rotation / Scale Push / Pop).  The position stack of the
// sequential semantics is resolved here too: every primitive knows the chain of transforms in force (outermost first),
// and the instruction stream carries "set the evaluation point" instructions — q = step(p) for the first step of a
// chain, q = step(q) for the following ones — emitted only when the chain changes (a chain that extends the previous
// one only emits the new steps).  Pops emit nothing, except ScalePop: acc = acc * s.  The operations a ray executes on
// its way to a primitive are therefore exactly those of the push commands enclosing it, in order: same bits as the
// oracle's position stack.  Programs with such commands are `extended` and run on the kernels' EXT instantiations;
// all other programs decode exactly as before and run on the unchanged ones.
#pragma once
#ifdef __CUDACC_RTC__  // runtime compilation of specialised kernels (rm_jit.cu): no host headers there
typedef unsigned char uint8_t;
typedef unsigned int uint32_t;
typedef unsigned long long uint64_t;
#else
#include <stdint.h>

#include <vector>
#endif

#ifndef RM_STACK_MAX
#define RM_STACK_MAX 32u  // value_stack_max_size, ray_marching.wgsl:173 (same value as rm_b200.h)
#endif

namespace rm {

// combine modes (low 3 bits of RmInsn::op)
enum : uint32_t {
    CB_FIRST = 0,   // acc = v                       (stack was empty)
    CB_PUSH = 1,    // spill[slot] = acc; acc = v
    CB_UNION = 2,   // acc = min(a, b)               wgsl:242-246
    CB_SUB = 3,     // acc = max(a, -b)              wgsl:248-252
    CB_INTER = 4,   // acc = max(a, b)               X1
    CB_SMOOTH = 5,  // acc = smooth-min(a, b; k)     X2
};
// operand source (bits 3..4 of RmInsn::op)
enum : uint32_t {
    SRC_POP = 0,     // b = acc, a = spill[slot]     (binary op whose right operand is a subtree)
    SRC_SPHERE = 1,  // b = sphere(p), a = acc       wgsl:229-233
    SRC_BOX = 2,     // b = box(p),    a = acc       wgsl:235-240
    SRC_PLANE = 3,   // b = plane(p),  a = acc       N4 (extended programs only)
};
// extended instructions that are not value producers (RmInsn::op; FLAG_XF / FLAG_ACCMUL set)
enum : uint32_t {
    XOP_TRANSLATE = 32,  // q = base - (cx,cy,cz)
    XOP_ROTATE = 33,     // q = transpose(M) * base, M row-major in cx,cy,cz,r0,r1,r2,k,inv_k,c4k
    XOP_SCALE = 34,      // q = base / cx   (fast mode: base * cy, cy = 1/s)
    XOP_RESET = 35,      // q = p           (the chain in force became empty)
    XOP_ACCMUL = 36,     // acc = acc * cx  (ScalePop)
};
// RmInsn::flags: one-hot description of the fused op, so that the interpreter dispatches with
// single-bit tests (LOP3 -> predicate, branch) in frequency order instead of a jump table.
enum : uint32_t {
    FLAG_K_POW2 = 1u << 0,  // 1/k is exact: multiply instead of divide in exact mode
    FLAG_BOX = 1u << 1,
    FLAG_POP = 1u << 2,
    FLAG_SMOOTH = 1u << 3,
    FLAG_UNION = 1u << 4,
    FLAG_SUB = 1u << 5,
    FLAG_INTER = 1u << 6,
    FLAG_PUSH = 1u << 7,
    FLAG_FIRST = 1u << 8,
    // extended programs only (N4)
    FLAG_PLANE = 1u << 9,
    FLAG_XF = 1u << 10,      // XOP_TRANSLATE / ROTATE / SCALE / RESET; slot bit 0: base = p (first step of a chain), else base = q
    FLAG_ACCMUL = 1u << 11,  // XOP_ACCMUL
    FLAG_SAT_OK = 1u << 12,  // FAST form only: the smooth union's k*s <= 1, so max(k*s - s|a-b|, 0) is a saturate (smooth_constants_fast)
};

// RmInsn::op is a DENSE fused opcode 0..15 so that the interpreter dispatches through a single
// jump table:  0..3 = POP x {UNION,SUB,INTER,SMOOTH};  4..9 = SPHERE x {FIRST..SMOOTH};
// 10..15 = BOX x {FIRST..SMOOTH}; extended programs add 16..21 = PLANE x {FIRST..SMOOTH} and the XOP_* values.
#define RM_HD
#ifdef __CUDACC__
#undef RM_HD
#define RM_HD __host__ __device__
#endif
RM_HD constexpr uint32_t opcode(uint32_t src, uint32_t cb) {
    return src == SRC_POP ? cb - CB_UNION : (src == SRC_SPHERE ? 4u + cb : (src == SRC_BOX ? 10u + cb : 16u + cb));
}
RM_HD constexpr uint32_t opcode_src(uint32_t op) { return op < 4u ? SRC_POP : (op < 10u ? SRC_SPHERE : SRC_BOX); }
RM_HD constexpr uint32_t opcode_cb(uint32_t op) { return op < 4u ? op + CB_UNION : (op < 10u ? op - 4u : op - 10u); }
// the same for extended programs (value-producing instructions only: op < 22)
RM_HD constexpr uint32_t opcode_src_ext(uint32_t op) { return op < 16u ? opcode_src(op) : SRC_PLANE; }
RM_HD constexpr uint32_t opcode_cb_ext(uint32_t op) { return op < 16u ? opcode_cb(op) : op - 16u; }

// One decoded instruction = 3 x 16 bytes (three broadcast LDS.128).
struct alignas(16) RmInsn {
    float cx, cy, cz, r0;      // centre, sphere radius / box radius x
    float r1, r2, k, inv_k;    // box radius y,z; smooth-union k and 1/k.  FAST form of the program: inv_k holds k*s instead,
    float c4k;                 // and c4k holds s = sqrt(1/(4k)) (smooth_constants_fast); exact form: c4k = 1/(4k), unread
    uint32_t op;               // dense fused opcode, see opcode()
    uint32_t slot;             // spill slot read (SRC_POP) or written (CB_PUSH)
    uint32_t flags;
};
static_assert(sizeof(RmInsn) == 48, "RmInsn must be 48 bytes");

#ifndef __CUDACC_RTC__
struct Program {
    std::vector<RmInsn> insns;
    uint32_t cmd_count = 0;
    uint32_t max_depth = 0;   // value-stack depth of the original program (<= 32)
    uint32_t n_spill = 0;     // spill slots used (= max_depth - 1, or 0)
    uint32_t n_sphere = 0, n_box = 0, n_union = 0, n_sub = 0, n_inter = 0, n_smooth = 0;
    // N4: any Plane / transform command present -> the kernels' EXT instantiations run it
    bool extended = false;
    uint32_t n_plane = 0, n_translate = 0, n_rotate = 0, n_scale = 0;
    // algorithmic flops of one map_scene evaluation (SURVEY.md §8(a) FLOP table; N4, counted the same literal way on
    // the sequential semantics: plane 3 mul + 3 add, translation 3 sub, rotation 9 mul + 6 add, scale 3 div + 1 mul)
    uint32_t flops_per_eval() const {
        return 10 * n_sphere + 22 * n_box + n_union + 2 * n_sub + n_inter + 10 * n_smooth + 6 * n_plane + 3 * n_translate +
               15 * n_rotate + 4 * n_scale;
    }
};

// Per-instruction constants of the smooth union (X2) — one definition for the interpreter's decoder and the
// specialised-kernel generator (rm_jit.cu), so that both evaluate the blend with the same numbers.
void smooth_constants(float k, float *inv_k, float *c4k, bool *k_pow2);
// Fast mode evaluates  min(a,b) - h*h*k/4  (h = max(k-|a-b|,0)/k)  as  min(a,b) - e*e  with
// e = max(ks - |a-b|*s, 0),  s = sqrt(1/(4k)),  ks = k*s:  FADD, FFMA, 2 FMNMX, FFMA — one instruction fewer than
// min(a,b) - (max(k-|a-b|,0))^2 * (1/(4k)).  For k a power of four (k = 0.25: s = 1) the two are the same numbers.
// When k*s <= 1 (k <= 4) the clamp at 0 can be the FFMA's own .SAT modifier: k*s - s|a-b| never exceeds k*s, so the clamp
// at 1 that comes with it never acts and the value is the same bit for bit — 4 instructions (FADD, FFMA.SAT, FMNMX, FFMA).
// The compiled kernels use it (the choice is part of the structure signature), and so does the 8-ray interpreter
// (FLAG_SAT_OK on the decoded instruction).
void smooth_constants_fast(float k, float *ks, float *s);
inline bool smooth_fast_sat_ok(float ks) { return ks <= 1.0f; }

// Returns an RM_* code (rm_b200.h).  `out` may be null for validation only.
// A prepared program is decoded twice (rm_abi.cu): the EXACT form keeps the sequential operation order and roundings of
// the oracle; the FAST form (what RM_MODE_FAST paints) differs in two ways, both algebraically equal to it:
//   fast_form:        smooth unions carry (k*s, s) in place of (1/k, 1/(4k)) — see smooth_constants_fast;
//   collapse_chains:  (extended programs) instead of one instruction per transform step, each primitive's chain of
//                     groups is composed on the host (double precision) into one similarity q = A*p + b — A is one
//                     "q = transpose(M)*p" instruction, shared by consecutive primitives with the same A and absent for
//                     translation-only chains, and b moves into the primitive's centre / the plane's offset.
//                     Measured on B200: S3 through the EXT interpreter 20.74 -> 16.76 ms (profiles/r2_s1_n4_collapse_timing.jsonl).
int decode_program(uint32_t cmd_count, const uint32_t *words, uint32_t n_words, uint32_t max_words,
                   Program *out, bool fast_form = false, bool collapse_chains = false);
#endif  // !__CUDACC_RTC__

}  // namespace rm
