// rm_program.cpp — see rm_program.h.  Pure host code, no CUDA.
#include "rm_program.h"

#include <cmath>
#include <algorithm>
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
    float p[9];
    uint32_t bits0;  // first parameter word as sent (ScalePop must repeat ScalePush's s bit for bit)
};
inline bool is_binary(uint32_t t) {
    return t == RM_OP_UNION || t == RM_OP_SUBTRACTION || t == RM_OP_INTERSECTION || t == RM_OP_SMOOTH_UNION;
}
inline bool is_primitive(uint32_t t) { return t == RM_OP_SPHERE || t == RM_OP_BOX || t == RM_OP_PLANE; }
inline bool is_xf_push(uint32_t t) { return t == RM_OP_TRANSLATION_PUSH || t == RM_OP_ROTATION_PUSH || t == RM_OP_SCALE_PUSH; }
inline bool is_xf_pop(uint32_t t) { return t == RM_OP_TRANSLATION_POP || t == RM_OP_ROTATION_POP || t == RM_OP_SCALE_POP; }
inline uint32_t combine_of(uint32_t t) {
    switch (t) {
        case RM_OP_UNION: return CB_UNION;
        case RM_OP_SUBTRACTION: return CB_SUB;
        case RM_OP_INTERSECTION: return CB_INTER;
        default: return CB_SMOOTH;
    }
}
inline void set_k(RmInsn &in, float k, bool fast_form) {
    in.k = k;
    bool pow2 = false;
    smooth_constants(k, &in.inv_k, &in.c4k, &pow2);
    if (pow2) in.flags |= FLAG_K_POW2;
    if (fast_form) {
        smooth_constants_fast(k, &in.inv_k, &in.c4k);  // (k*s, s) in the slots of (1/k, 1/(4k))
        if (smooth_fast_sat_ok(in.inv_k)) in.flags |= FLAG_SAT_OK;
    }
}
}  // namespace

void smooth_constants_fast(float k, float *ks, float *s) {
    const double sd = std::sqrt(0.25 / (double)k);
    *s = (float)sd;
    *ks = (float)((double)k * sd);
}

void smooth_constants(float k, float *inv_k, float *c4k, bool *k_pow2) {
    *inv_k = 1.0f / k;
    *c4k = 0.25f / k;  // fast mode folds h*h*k*0.25 = e*e/(4k) into one constant
    int e = 0;
    const float m = std::frexp(k, &e);
    // 1/k and x*(1/k) == x/k bit for bit when k is a normal power of two with a normal reciprocal
    *k_pow2 = m == 0.5f && std::fpclassify(k) == FP_NORMAL && std::fpclassify(*inv_k) == FP_NORMAL;
}

namespace {

void finish_flags(RmInsn &in) {
    const uint32_t src = opcode_src_ext(in.op), cb = opcode_cb_ext(in.op);
    static const uint32_t cb_flag[6] = {FLAG_FIRST, FLAG_PUSH, FLAG_UNION, FLAG_SUB, FLAG_INTER, FLAG_SMOOTH};
    in.flags |= cb_flag[cb] | (src == SRC_BOX ? FLAG_BOX : 0u) | (src == SRC_POP ? FLAG_POP : 0u) | (src == SRC_PLANE ? FLAG_PLANE : 0u);
}

// Programs with N4 commands (see rm_program.h).  `cmds` is the parsed stream; counts and `extended` are set in `prog`.
// The similarity q = A*p + b that a chain of pushes (outermost first) amounts to, in double precision.
struct Frame {
    double A[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};  // row-major
    double b[3] = {0, 0, 0};
    bool linear_is_identity() const {
        static const double I[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
        return std::equal(A, A + 9, I);
    }
};
Frame compose_chain(const std::vector<Cmd> &cmds, const std::vector<uint32_t> &chain) {
    Frame f;
    for (uint32_t idx : chain) {
        const Cmd &p = cmds[idx];
        if (p.type == RM_OP_TRANSLATION_PUSH) {  // q - t
            for (int i = 0; i < 3; ++i) f.b[i] -= (double)p.p[i];
        } else if (p.type == RM_OP_ROTATION_PUSH) {  // transpose(M) * q, M row-major in p.p
            double A2[9], b2[3];
            for (int i = 0; i < 3; ++i) {
                for (int j = 0; j < 3; ++j) A2[3 * i + j] = (double)p.p[i] * f.A[j] + (double)p.p[3 + i] * f.A[3 + j] + (double)p.p[6 + i] * f.A[6 + j];
                b2[i] = (double)p.p[i] * f.b[0] + (double)p.p[3 + i] * f.b[1] + (double)p.p[6 + i] * f.b[2];
            }
            std::copy(A2, A2 + 9, f.A);
            std::copy(b2, b2 + 3, f.b);
        } else {  // q / s
            for (double &x : f.A) x /= (double)p.p[0];
            for (double &x : f.b) x /= (double)p.p[0];
        }
    }
    return f;
}

// collapse_chains (fast arithmetic only; exact mode has to keep the sequential roundings): each primitive's chain is
// composed on the host into one similarity q = A*p + b; A becomes ONE "q = transpose(M)*p" instruction (none when the
// chain only translates), b is folded into the primitive's centre (sphere, box) or offset (plane).
int lower_extended(const std::vector<Cmd> &cmds, Program &prog, Program *out, bool fast_form, bool collapse_chains) {
    const uint32_t n = (uint32_t)cmds.size();
    // pass 2a: value-stack and group discipline, command by command as the sequential semantics would run
    // (same rules as oracle/rm_oracle.c validate_program)
    struct Group { uint32_t type, depth, s_bits; };
    std::vector<Group> groups;
    uint32_t depth = 0, max_depth = 0;
    for (uint32_t i = 0; i < n; ++i) {
        const Cmd &c = cmds[i];
        if (is_xf_push(c.type)) {
            if (groups.size() >= RM_POS_STACK_MAX) return RM_E_STACK_OVERFLOW;
            groups.push_back(Group{c.type, depth, c.bits0});
        } else if (is_xf_pop(c.type)) {
            // closes the innermost group: same kind, and exactly one value was produced inside it
            if (groups.empty() || groups.back().type + 1u != c.type || depth != groups.back().depth + 1u) return RM_E_STACK_UNBALANCED;
            if (c.type == RM_OP_SCALE_POP && c.bits0 != groups.back().s_bits) return RM_E_BAD_PARAM;
            groups.pop_back();
        } else if (is_binary(c.type)) {
            // an operator inside a group may only consume values produced inside it
            if (depth - (groups.empty() ? 0u : groups.back().depth) < 2u) return RM_E_STACK_UNBALANCED;
            depth -= 1;
        } else {
            if (depth + 1 > RM_STACK_MAX) return RM_E_STACK_OVERFLOW;
            depth += 1;
            if (depth > max_depth) max_depth = depth;
        }
    }
    if (depth < 1 || !groups.empty()) return RM_E_STACK_UNBALANCED;

    // pass 2b: resolve the position stack.  Value commands keep their order; every primitive gets the chain of push
    // commands in force (indices into cmds, outermost first); T/R pops vanish, ScalePop becomes acc *= s.
    struct VCmd { uint32_t cmd; bool accmul; std::vector<uint32_t> chain; };  // collapse_chains: chain is composed at emission
    std::vector<VCmd> v;
    std::vector<uint32_t> open;
    for (uint32_t i = 0; i < n; ++i) {
        const Cmd &c = cmds[i];
        if (is_xf_push(c.type)) open.push_back(i);
        else if (is_xf_pop(c.type)) {
            open.pop_back();
            if (c.type == RM_OP_SCALE_POP) v.push_back(VCmd{i, true, {}});
        } else v.push_back(VCmd{i, false, is_primitive(c.type) ? open : std::vector<uint32_t>()});
    }

    // pass 2c: emission — the reference-set algorithm (top of stack in the accumulator, primitive + operator fused)
    // plus "set the evaluation point" instructions whenever the chain in force changes
    std::vector<uint32_t> cur;  // chain whose result q holds (empty: q == p, the state at the start of every evaluation)
    auto emit_chain = [&](const std::vector<uint32_t> &chain) {
        if (chain == cur) return;
        if (chain.empty()) {
            RmInsn in{};
            in.op = XOP_RESET; in.flags = FLAG_XF; in.slot = 1;
            prog.insns.push_back(in);
        } else {
            size_t start = 0;  // a chain that extends the current one continues from q
            if (!cur.empty() && cur.size() < chain.size() && std::equal(cur.begin(), cur.end(), chain.begin())) start = cur.size();
            for (size_t k = start; k < chain.size(); ++k) {
                const Cmd &p = cmds[chain[k]];
                RmInsn in{};
                in.flags = FLAG_XF;
                in.slot = k == 0 ? 1u : 0u;  // bit 0: base = p
                if (p.type == RM_OP_TRANSLATION_PUSH) {
                    in.op = XOP_TRANSLATE; in.cx = p.p[0]; in.cy = p.p[1]; in.cz = p.p[2];
                } else if (p.type == RM_OP_ROTATION_PUSH) {
                    in.op = XOP_ROTATE;
                    in.cx = p.p[0]; in.cy = p.p[1]; in.cz = p.p[2]; in.r0 = p.p[3]; in.r1 = p.p[4]; in.r2 = p.p[5];
                    in.k = p.p[6]; in.inv_k = p.p[7]; in.c4k = p.p[8];
                } else {
                    in.op = XOP_SCALE; in.cx = p.p[0]; in.cy = 1.0f / p.p[0];
                }
                prog.insns.push_back(in);
            }
        }
        cur = chain;
    };
    // collapse_chains: the linear part q currently holds (as the float matrix M = transpose(A) of its instruction)
    bool cur_identity = true;
    float cur_m[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    auto emit_linear = [&](const Frame &f) {
        float m[9];
        for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) m[3 * i + j] = (float)f.A[3 * j + i];
        const bool ident = f.linear_is_identity();
        if (ident ? cur_identity : (!cur_identity && std::memcmp(m, cur_m, sizeof m) == 0)) return;
        RmInsn in{};
        in.flags = FLAG_XF;
        in.slot = 1;  // always from the ray position
        if (ident) {
            in.op = XOP_RESET;
        } else {
            in.op = XOP_ROTATE;
            in.cx = m[0]; in.cy = m[1]; in.cz = m[2]; in.r0 = m[3]; in.r1 = m[4]; in.r2 = m[5]; in.k = m[6]; in.inv_k = m[7]; in.c4k = m[8];
        }
        prog.insns.push_back(in);
        cur_identity = ident;
        std::memcpy(cur_m, m, sizeof m);
    };
    depth = 0;
    const uint32_t nv = (uint32_t)v.size();
    for (uint32_t i = 0; i < nv; ++i) {
        const Cmd &c = cmds[v[i].cmd];
        RmInsn in{};
        if (v[i].accmul) {
            in.op = XOP_ACCMUL; in.flags = FLAG_ACCMUL; in.cx = c.p[0];
            prog.insns.push_back(in);
            continue;
        }
        if (!is_binary(c.type)) {
            in.cx = c.p[0]; in.cy = c.p[1]; in.cz = c.p[2]; in.r0 = c.p[3];
            if (collapse_chains) {
                const Frame f = compose_chain(cmds, v[i].chain);
                emit_linear(f);
                if (c.type == RM_OP_PLANE) {  // dot(A p + b, n) + h = dot(A p, n) + (dot(b, n) + h)
                    in.r0 = (float)((double)c.p[3] + f.b[0] * (double)c.p[0] + f.b[1] * (double)c.p[1] + f.b[2] * (double)c.p[2]);
                } else {  // |A p + b - c| = |A p - (c - b)|
                    in.cx = (float)((double)c.p[0] - f.b[0]); in.cy = (float)((double)c.p[1] - f.b[1]); in.cz = (float)((double)c.p[2] - f.b[2]);
                }
            } else {
                emit_chain(v[i].chain);
            }
            uint32_t src = c.type == RM_OP_SPHERE ? SRC_SPHERE : (c.type == RM_OP_BOX ? SRC_BOX : SRC_PLANE);
            if (c.type == RM_OP_BOX) { in.r1 = c.p[4]; in.r2 = c.p[5]; }
            if (depth >= 1 && i + 1 < nv && !v[i + 1].accmul && is_binary(cmds[v[i + 1].cmd].type)) {
                const Cmd &o = cmds[v[i + 1].cmd];  // fused: acc = op(acc, prim); T/R pops in between change nothing
                in.op = opcode(src, combine_of(o.type));
                if (o.type == RM_OP_SMOOTH_UNION) set_k(in, o.p[0], fast_form);
                ++i;
            } else if (depth == 0) {
                in.op = opcode(src, CB_FIRST);
                depth = 1;
            } else {
                in.op = opcode(src, CB_PUSH);
                in.slot = depth - 1;
                depth += 1;
            }
        } else {
            in.op = opcode(SRC_POP, combine_of(c.type));
            in.slot = depth - 2;
            if (c.type == RM_OP_SMOOTH_UNION) set_k(in, c.p[0], fast_form);
            depth -= 1;
        }
        finish_flags(in);
        prog.insns.push_back(in);
    }
    prog.cmd_count = n;
    prog.max_depth = max_depth;
    prog.n_spill = max_depth > 0 ? max_depth - 1 : 0;
    if (out) *out = std::move(prog);
    return RM_OK;
}

}  // namespace

int decode_program(uint32_t cmd_count, const uint32_t *words, uint32_t n_words, uint32_t max_words,
                   Program *out, bool fast_form, bool collapse_chains) {
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
            case RM_OP_PLANE: np = 4; prog.n_plane++; prog.extended = true; break;
            case RM_OP_TRANSLATION_PUSH: np = 3; prog.n_translate++; prog.extended = true; break;
            case RM_OP_ROTATION_PUSH: np = 9; prog.n_rotate++; prog.extended = true; break;
            case RM_OP_SCALE_PUSH: np = 1; prog.n_scale++; prog.extended = true; break;
            case RM_OP_TRANSLATION_POP: case RM_OP_ROTATION_POP: prog.extended = true; break;
            case RM_OP_SCALE_POP: np = 1; prog.extended = true; break;
            default: return RM_E_BAD_OPCODE;
        }
        if (ptr + np > n_words) return RM_E_STACK_UNBALANCED;
        for (uint32_t j = 0; j < np; ++j) c.p[j] = as_f32(words[ptr + j]);
        if (np) c.bits0 = words[ptr];
        ptr += np;
        if ((c.type == RM_OP_SMOOTH_UNION || c.type == RM_OP_SCALE_PUSH) && !(c.p[0] > 0.0f && std::isfinite(c.p[0])))
            return RM_E_BAD_PARAM;
        // A NaN parameter makes every distance NaN in the shader; the kernels' min / max (fminf / fmaxf drop a NaN
        // operand) and the literal `(b < a) ? b : a` of the restatement then part ways, so it is refused here and by
        // the oracle's validator alike.  Infinite values are accepted (a box radius of +inf is a slab).
        for (uint32_t j = 0; j < np; ++j) if (c.p[j] != c.p[j]) return RM_E_BAD_PARAM;
        cmds.push_back(c);
    }
    if (prog.extended) return lower_extended(cmds, prog, out, fast_form, collapse_chains);

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
                if (o.type == RM_OP_SMOOTH_UNION) set_k(in, o.p[0], fast_form);
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
            if (c.type == RM_OP_SMOOTH_UNION) set_k(in, c.p[0], fast_form);
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
