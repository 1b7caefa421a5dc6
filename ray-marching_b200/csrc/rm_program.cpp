// rm_program.cpp — see rm_program.h.  Pure host code, no CUDA.
#include "rm_program.h"

#include <cmath>
#include <cstring>

#include "../../include/rm_b200.h"

namespace rm {

namespace {
inline float as_f32(uint32_t u) {
    float f;
    std::memcpy(&f, &u, 4);
    return f;
}
struct Cmd {
    uint32_t type;
    float p[6];
};
inline bool is_binary(uint32_t t) {
    return t == RM_OP_UNION || t == RM_OP_SUBTRACTION || t == RM_OP_INTERSECTION || t == RM_OP_SMOOTH_UNION;
}
inline uint32_t combine_of(uint32_t t) {
    switch (t) {
        case RM_OP_UNION: return CB_UNION;
        case RM_OP_SUBTRACTION: return CB_SUB;
        case RM_OP_INTERSECTION: return CB_INTER;
        default: return CB_SMOOTH;
    }
}
inline void set_k(RmInsn &in, float k) {
    in.k = k;
    bool pow2 = false;
    smooth_constants(k, &in.inv_k, &in.c4k, &pow2);
    if (pow2) in.flags |= FLAG_K_POW2;
}
}  // namespace

void smooth_constants(float k, float *inv_k, float *c4k, bool *k_pow2) {
    *inv_k = 1.0f / k;
    *c4k = 0.25f / k;  // fast mode folds h*h*k*0.25 = e*e/(4k) into one constant
    int e = 0;
    const float m = std::frexp(k, &e);
    // 1/k and x*(1/k) == x/k bit for bit when k is a normal power of two with a normal reciprocal
    *k_pow2 = m == 0.5f && std::fpclassify(k) == FP_NORMAL && std::fpclassify(*inv_k) == FP_NORMAL;
}

int decode_program(uint32_t cmd_count, const uint32_t *words, uint32_t n_words, uint32_t max_words,
                   Program *out) {
    if (out) *out = Program();
    if (cmd_count == 0) return RM_OK;  // empty scene: map_scene == max_dist (wgsl:188-191)
    if (!words) return RM_E_INVALID_ARG;
    if (n_words > max_words) return RM_E_PROGRAM_TOO_LARGE;

    // pass 1: parse the word stream exactly as csg_pop_* would walk it (wgsl:154-170)
    std::vector<Cmd> cmds;
    cmds.reserve(cmd_count);
    Program prog;
    uint32_t ptr = 0;
    for (uint32_t i = 0; i < cmd_count; ++i) {
        if (ptr >= n_words) return RM_E_STACK_UNBALANCED;
        Cmd c{};
        c.type = words[ptr++];
        uint32_t np = 0;
        switch (c.type) {
            case RM_OP_SPHERE: np = 4; prog.n_sphere++; break;
            case RM_OP_BOX: np = 6; prog.n_box++; break;
            case RM_OP_UNION: prog.n_union++; break;
            case RM_OP_SUBTRACTION: prog.n_sub++; break;
            case RM_OP_INTERSECTION: prog.n_inter++; break;
            case RM_OP_SMOOTH_UNION: np = 1; prog.n_smooth++; break;
            default: return RM_E_BAD_OPCODE;
        }
        if (ptr + np > n_words) return RM_E_STACK_UNBALANCED;
        for (uint32_t j = 0; j < np; ++j) c.p[j] = as_f32(words[ptr + j]);
        ptr += np;
        if (c.type == RM_OP_SMOOTH_UNION && !(c.p[0] > 0.0f && std::isfinite(c.p[0]))) return RM_E_BAD_PARAM;
        cmds.push_back(c);
    }

    // pass 2: stack discipline + emission with top-of-stack caching and prim/op fusion
    uint32_t depth = 0, max_depth = 0;
    for (uint32_t i = 0; i < cmd_count; ++i) {
        const Cmd &c = cmds[i];
        RmInsn in{};
        if (!is_binary(c.type)) {
            if (depth + 1 > RM_STACK_MAX) return RM_E_STACK_OVERFLOW;
            if (depth + 1 > max_depth) max_depth = depth + 1;
            in.cx = c.p[0]; in.cy = c.p[1]; in.cz = c.p[2]; in.r0 = c.p[3];
            uint32_t src = SRC_SPHERE;
            if (c.type == RM_OP_BOX) { src = SRC_BOX; in.r1 = c.p[4]; in.r2 = c.p[5]; }
            if (depth >= 1 && i + 1 < cmd_count && is_binary(cmds[i + 1].type)) {
                const Cmd &o = cmds[i + 1];  // fused: acc = op(acc, prim)
                in.op = opcode(src, combine_of(o.type));
                if (o.type == RM_OP_SMOOTH_UNION) set_k(in, o.p[0]);
                ++i;  // the operator is consumed; depth unchanged (push 1, pop 2, push 1)
            } else if (depth == 0) {
                in.op = opcode(src, CB_FIRST);
                depth = 1;
            } else {
                in.op = opcode(src, CB_PUSH);
                in.slot = depth - 1;
                depth += 1;
            }
        } else {
            if (depth < 2) return RM_E_STACK_UNBALANCED;
            in.op = opcode(SRC_POP, combine_of(c.type));
            in.slot = depth - 2;
            if (c.type == RM_OP_SMOOTH_UNION) set_k(in, c.p[0]);
            depth -= 1;
        }
        {
            const uint32_t src = opcode_src(in.op), cb = opcode_cb(in.op);
            static const uint32_t cb_flag[6] = {FLAG_FIRST, FLAG_PUSH, FLAG_UNION, FLAG_SUB, FLAG_INTER, FLAG_SMOOTH};
            in.flags |= cb_flag[cb] | (src == SRC_BOX ? FLAG_BOX : 0u) | (src == SRC_POP ? FLAG_POP : 0u);
        }
        prog.insns.push_back(in);
    }
    if (depth < 1) return RM_E_STACK_UNBALANCED;
    prog.cmd_count = cmd_count;
    prog.max_depth = max_depth;
    prog.n_spill = max_depth > 0 ? max_depth - 1 : 0;
    if (out) *out = std::move(prog);
    return RM_OK;
}

}  // namespace rm
