// rm_jit.cu — see rm_jit.h.  Emits map_scene of one command stream as straight-line CUDA, compiles it with NVRTC
// (loaded with dlopen: the library has no link-time dependency on it) for sm_100a and loads the cubin through the
// runtime's library API.  Everything device-side comes from the same headers the ahead-of-time kernels use; they are
// embedded in this object as strings (rm_jit_headers.inc, written by the Makefile) and handed to NVRTC by name.
#include "rm_jit.h"

#include <dlfcn.h>

#include <chrono>
#include <condition_variable>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <vector>

#include "../../include/rm_b200.h"
#include "rm_direct_body.cuh"
#include "rm_program.h"

namespace rm {

namespace {

struct EmbeddedHeader {
    const char *name;
    const char *text;
};
const EmbeddedHeader kHeaders[] = {
#include "rm_jit_headers.inc"
};
constexpr int kNumHeaders = (int)(sizeof(kHeaders) / sizeof(kHeaders[0]));

// ---- code-size model (SASS instructions per ray of one map_scene evaluation, fast mode; counted in the generated
// code of the S0/S1/S2 scenes: sphere 3 FADD + FMUL + 2 FFMA + MUFU + FADD, box 6 FADD + 5 FMNMX + FMUL + 2 FFMA + MUFU
// + FADD, smooth union 2 FADD + 2 FMNMX + FMUL + FFMA) and the instruction-cache budget it is checked against ----
constexpr uint32_t kInsnSphere = 8, kInsnBox = 17, kInsnMinMax = 1, kInsnSmooth = 5;
// N4 commands, counted from their device functions (not yet measured): plane 3 FFMA + FMUL-free, translation 3 FADD,
// rotation 3 FMUL + 6 FFMA, scale 3 FMUL (push) + 1 FMUL (pop)
constexpr uint32_t kInsnPlane = 3, kInsnTranslate = 3, kInsnRotate = 9, kInsnScalePush = 3, kInsnScalePop = 1;
constexpr uint32_t kBytesPerInsn = 16;
constexpr size_t kMaxParamBytes = 3584;  // SpecParams + FrameParams + pointers stay inside the classic 4-KB kernel-argument space
constexpr uint32_t kLoopBudget4 = 24u << 10;  // 4 rays per thread while the march loop's map code stays under this
constexpr uint32_t kLoopBudget2 = 40u << 10;  // 2 rays per thread up to here; beyond it the interpreter is the better kernel

inline float as_f32(uint32_t u) {
    float f;
    std::memcpy(&f, &u, 4);
    return f;
}
inline uint32_t as_u32(float f) {
    uint32_t u;
    std::memcpy(&u, &f, 4);
    return u;
}
std::string lit(uint32_t bits) {  // a float constant by its bits: no decimal round trip
    char b[40];
    std::snprintf(b, sizeof b, "__uint_as_float(0x%08xu)", bits);
    return b;
}
std::string var(uint32_t i) { return "v" + std::to_string(i); }

int env_int(const char *name) {
    const char *s = std::getenv(name);
    return s ? std::atoi(s) : 0;
}

// ---- NVRTC through dlopen ----------------------------------------------------------------------
typedef struct _nvrtcProgram *nvrtcProgram;
struct Nvrtc {
    void *h = nullptr;
    int (*CreateProgram)(nvrtcProgram *, const char *, const char *, int, const char *const *, const char *const *) = nullptr;
    int (*DestroyProgram)(nvrtcProgram *) = nullptr;
    int (*CompileProgram)(nvrtcProgram, int, const char *const *) = nullptr;
    int (*GetProgramLogSize)(nvrtcProgram, size_t *) = nullptr;
    int (*GetProgramLog)(nvrtcProgram, char *) = nullptr;
    int (*GetCUBINSize)(nvrtcProgram, size_t *) = nullptr;
    int (*GetCUBIN)(nvrtcProgram, char *) = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
    std::string why;  // load failure
    int version = 0;  // major * 1000 + minor of the library in use
    std::string path;
};

const Nvrtc &nvrtc() {
    static const Nvrtc n = [] {
        Nvrtc r;
        std::vector<std::string> names;
        if (const char *p = std::getenv("RM_NVRTC_PATH")) names.push_back(p);  // exact file; nothing else is tried
        else {
            if (const char *home = std::getenv("CUDA_HOME")) names.push_back(std::string(home) + "/lib64/libnvrtc.so.12");
            names.insert(names.end(), {"/usr/local/cuda/lib64/libnvrtc.so.12", "/usr/local/cuda/lib64/libnvrtc.so", "libnvrtc.so.12", "libnvrtc.so"});
        }
        // Several NVRTC builds can be around — a process that imported torch already holds torch's bundled libnvrtc.so.12
        // (12.8 in this image), and a bare soname resolves to that one.  The NEWEST loadable candidate is used: the 12.8
        // compiler builds the packed-FP32 scene code 24 % slower than 12.9 (cfg3 33.4 against 26.9 ms, measured).
        int best_ver = -1;
        for (const auto &nm : names) {
            void *hnd = dlopen(nm.c_str(), RTLD_NOW | RTLD_LOCAL);
            if (!hnd) {
                const char *de = dlerror();
                r.why += nm + ": " + (de ? de : "not loadable") + "; ";
                continue;
            }
            int major = 0, minor = 0;
            auto ver = (int (*)(int *, int *))dlsym(hnd, "nvrtcVersion");
            const int v = (ver && ver(&major, &minor) == 0) ? major * 1000 + minor : 0;
            if (v > best_ver) {
                if (r.h) dlclose(r.h);
                r.h = hnd;
                best_ver = v;
                r.version = v;
                r.path = nm;
            } else dlclose(hnd);
        }
        if (!r.h) return r;
        bool ok = true;
        auto sym = [&](const char *s) {
            void *p = dlsym(r.h, s);
            if (!p) { ok = false; r.why += std::string("missing symbol ") + s + "; "; }
            return p;
        };
        r.CreateProgram = (decltype(r.CreateProgram))sym("nvrtcCreateProgram");
        r.DestroyProgram = (decltype(r.DestroyProgram))sym("nvrtcDestroyProgram");
        r.CompileProgram = (decltype(r.CompileProgram))sym("nvrtcCompileProgram");
        r.GetProgramLogSize = (decltype(r.GetProgramLogSize))sym("nvrtcGetProgramLogSize");
        r.GetProgramLog = (decltype(r.GetProgramLog))sym("nvrtcGetProgramLog");
        r.GetCUBINSize = (decltype(r.GetCUBINSize))sym("nvrtcGetCUBINSize");
        r.GetCUBIN = (decltype(r.GetCUBIN))sym("nvrtcGetCUBIN");
        r.GetErrorString = (decltype(r.GetErrorString))sym("nvrtcGetErrorString");
        if (!ok) { dlclose(r.h); r.h = nullptr; }
        return r;
    }();
    return n;
}

// Where a command parameter of the generated code comes from.  Default: the kernel's SpecParams argument, or for deep
// programs the module's __constant__ copy of it (constant bank either way, read through uniform registers: one LDCU per
// 1-4 values on sm_100 — arithmetic takes UR operands, not c[bank][off] ones) — the code then depends on the program's
// STRUCTURE only and one cubin serves every edit of a radius, a centre or a k.  Exception:
// a parameter that is exactly +0 or -0 stays a literal (and its zero-ness is part of the structure signature), so that
// `p - 0` and friends still fold; an editor drag leaves 0 once, not per frame.  The LITERAL tier emits every parameter
// as an immediate by its bits and is keyed by the whole word stream: sm_100 arithmetic takes immediates for free, but a
// kernel-argument value has to be loaded into a uniform register first (LDCU: one issue slot per 1-4 values; S2's march
// loop is 1165 instructions with arguments against 1105 with immediates), so words that stay put are worth a second
// compile (rm_abi_paint.inc current_spec promotes them in the background).
struct ParamSink {
    bool literals;
    std::vector<float> *params;      // may be null (source only)
    std::vector<uint32_t> *sig;      // may be null
    uint32_t n = 0;
    // The parameters of one command are read together: start them on a 16-byte (3 or more values) or 8-byte (2 values)
    // boundary of the block, so that they arrive in the uniform registers as one LDCU.128 / LDCU.64 instead of one load
    // each (sm_100 arithmetic takes uniform-register operands, not c[bank][offset] ones: every value costs a load slot)
    void group(const uint32_t *bits, uint32_t count) {
        if (literals) return;
        uint32_t nz = 0;
        for (uint32_t j = 0; j < count; ++j) nz += (bits[j] & 0x7fffffffu) != 0u;
        const uint32_t al = nz >= 3 ? 4u : (nz == 2 ? 2u : 1u);
        while (n % al) {
            if (params) params->push_back(0.0f);
            ++n;
        }
    }
    std::string operator()(uint32_t bits) {
        if (literals) return lit(bits);
        if ((bits & 0x7fffffffu) == 0u) {           // +-0: structural
            if (sig) sig->push_back(0xF0000000u | (bits >> 31));
            return bits ? "(-0.0f)" : "0.0f";
        }
        if (sig) sig->push_back(0xF0000002u);
        if (params) params->push_back(as_f32(bits));
        return "@P@.f[" + std::to_string(n++) + "]";  // @P@: the kernel argument or the module's __constant__ copy, decided below
    }
};

// ---- process-wide cache ------------------------------------------------------------------------
// Keyed by the structure signature of spec_generate_source (+ mode, variant); LRU-bounded.  Entries are shared_ptrs:
// a context keeps the kernels of its current program alive, eviction only drops the cache's reference, and the cubin is
// unloaded (SpecKernel's destructor) when the last holder lets go.
struct Key {
    std::vector<uint32_t> sig;
    bool exact;
    int variant;
    bool operator<(const Key &o) const {
        if (exact != o.exact) return exact < o.exact;
        if (variant != o.variant) return variant < o.variant;
        return sig < o.sig;
    }
};
struct Entry {
    int rc = RM_OK;
    std::string err;
    std::shared_ptr<SpecKernel> k;
    uint64_t last_use = 0;
    bool ready = false;  // false while its compile is in flight (the mutex is not held meanwhile)
};
size_t g_cache_cap = 512;
uint64_t g_tick = 0, g_evictions = 0;
std::mutex g_mu;
std::condition_variable g_cv;
std::map<Key, Entry> g_cache;

// with g_mu held: make room for one more entry by dropping the least recently used one nobody else holds
bool evict_one() {
    auto victim = g_cache.end();
    for (auto it = g_cache.begin(); it != g_cache.end(); ++it) {
        if (!it->second.ready) continue;                               // being compiled
        if (it->second.k && it->second.k.use_count() > 1) continue;  // painted by some context right now
        if (victim == g_cache.end() || it->second.last_use < victim->second.last_use) victim = it;
    }
    if (victim == g_cache.end()) return false;
    g_cache.erase(victim);  // ~SpecKernel unloads the cubin
    ++g_evictions;
    return true;
}

}  // namespace

SpecKernel::~SpecKernel() {
    if (!lib) return;
    // kernels of this cubin may still be running on the devices that launched it
    int cur = 0;
    cudaGetDevice(&cur);
    const uint64_t m = dev_mask.load();
    for (int d = 0; d < 64; ++d)
        if (m & (1ull << d)) { cudaSetDevice(d); cudaDeviceSynchronize(); }
    for (auto &kv : cb)
        if (kv.second.last_launch) { cudaSetDevice(kv.first); cudaEventDestroy(kv.second.last_launch); }
    cudaSetDevice(cur);
    cudaLibraryUnload(lib);
}

void spec_cache_set_capacity(size_t n) {
    std::lock_guard<std::mutex> lock(g_mu);
    g_cache_cap = n ? n : 1;
    while (g_cache.size() > g_cache_cap && evict_one()) {}
}
void spec_cache_stats(size_t *entries, size_t *capacity, uint64_t *evictions) {
    std::lock_guard<std::mutex> lock(g_mu);
    if (entries) *entries = g_cache.size();
    if (capacity) *capacity = g_cache_cap;
    if (evictions) *evictions = g_evictions;
}

int spec_generate_source(uint32_t cmd_count, const uint32_t *words, uint32_t n_words, bool exact, int variant, int R,
                         std::string *src, int *r_out, int *minb_out, uint32_t *est_out, std::string *err,
                         std::vector<float> *params, std::vector<uint32_t> *sig, bool literals, bool *constbank_out) {
    // walk the postfix stream (builder.rs:26-62) with a stack of SSA value numbers: command i defines v<i>.
    // The stream has been validated by decode_program (rm_prepare / rm_spec_source); the checks here only keep the walk
    // itself in bounds and accept exactly what the decoder accepts.
    ParamSink par{literals, params, sig};
    if (params) params->clear();
    if (sig) {
        sig->clear();
        if (literals) { sig->push_back(0xFFFFFFFFu); sig->push_back(cmd_count); sig->insert(sig->end(), words, words + n_words); }
    }
    auto sig_op = [&](uint32_t op) { if (sig && !literals) sig->push_back(op); };
    std::string body;
    // the same scene code on PAIRS of rays (f2_t, rm_device.cuh: FADD2 / FMUL2 / FFMA2): fast mode, reference-set programs
    std::string body2;
    bool packable = !exact;
    auto wvar = [](uint32_t i) { return "w" + std::to_string(i); };
    std::vector<uint32_t> stack;
    // N4: the evaluation point is an SSA value too — a Push defines q<i>x/y/z from the point in force, the Pop goes back
    // to the previous names.  The root point is the functor's argument (px, py, pz).
    struct Pos { std::string x, y, z; uint32_t push_op, s_bits; size_t depth; };
    std::vector<Pos> pos = {Pos{"px", "py", "pz", 0, 0, 0}};
    uint32_t ptr = 0, est = 0;
    auto need = [&](uint32_t n) { return ptr + n <= n_words; };
    auto at = [&]() { return pos.back().x + "[r], " + pos.back().y + "[r], " + pos.back().z + "[r]"; };
    for (uint32_t i = 0; i < cmd_count; ++i) {
        if (!need(1)) { if (err) *err = "truncated command stream"; return RM_E_STACK_UNBALANCED; }
        const uint32_t op = words[ptr++];
        sig_op(op);
        std::string e, e2;
        if (op == RM_OP_PLANE || (op >= RM_OP_TRANSLATION_PUSH && op <= RM_OP_SCALE_POP)) packable = false;
        if (op == RM_OP_SPHERE) {
            if (!need(4)) { if (err) *err = "truncated sphere"; return RM_E_STACK_UNBALANCED; }
            par.group(words + ptr, 4);
            e = "sd_sphere<EX>(" + at();
            e2 = "sd_sphere2(PX[h], PY[h], PZ[h]";
            for (int j = 0; j < 4; ++j) { const std::string a_ = par(words[ptr + j]); e += ", " + a_; e2 += ", " + a_; }
            e += ")"; e2 += ")";
            ptr += 4; est += kInsnSphere;
            stack.push_back(i);
        } else if (op == RM_OP_BOX) {
            if (!need(6)) { if (err) *err = "truncated box"; return RM_E_STACK_UNBALANCED; }
            par.group(words + ptr, 6);
            e = "sd_box<EX>(" + at();
            e2 = "sd_box2(PX[h], PY[h], PZ[h]";
            for (int j = 0; j < 6; ++j) { const std::string a_ = par(words[ptr + j]); e += ", " + a_; e2 += ", " + a_; }
            e += ")"; e2 += ")";
            ptr += 6; est += kInsnBox;
            stack.push_back(i);
        } else if (op == RM_OP_PLANE) {
            if (!need(4)) { if (err) *err = "truncated plane"; return RM_E_STACK_UNBALANCED; }
            par.group(words + ptr, 4);
            e = "sd_plane<EX>(" + at();
            for (int j = 0; j < 4; ++j) e += ", " + par(words[ptr + j]);
            e += ")";
            ptr += 4; est += kInsnPlane;
            stack.push_back(i);
        } else if (op == RM_OP_TRANSLATION_PUSH || op == RM_OP_ROTATION_PUSH || op == RM_OP_SCALE_PUSH) {
            const uint32_t np = op == RM_OP_TRANSLATION_PUSH ? 3u : (op == RM_OP_ROTATION_PUSH ? 9u : 1u);
            if (!need(np)) { if (err) *err = "truncated transform"; return RM_E_STACK_UNBALANCED; }
            if (pos.size() > RM_POS_STACK_MAX) { if (err) *err = "transform groups nested deeper than 8"; return RM_E_STACK_OVERFLOW; }
            const std::string q = "q" + std::to_string(i);
            std::string call;
            if (op == RM_OP_TRANSLATION_PUSH) { call = "xf_translate<EX>(" + at(); est += kInsnTranslate; }
            else if (op == RM_OP_ROTATION_PUSH) { call = "xf_rotate<EX>(" + at(); est += kInsnRotate; }
            else {
                const float sc = as_f32(words[ptr]);
                if (!(sc > 0.0f) || !std::isfinite(sc)) { if (err) *err = "scale must be finite and > 0"; return RM_E_BAD_PARAM; }
                const uint32_t sb[2] = {words[ptr], as_u32(1.0f / sc)};
                par.group(sb, 2);
                call = "xf_scale<EX>(" + at() + ", " + par(sb[0]) + ", " + par(sb[1]);
                est += kInsnScalePush;
            }
            if (op != RM_OP_SCALE_PUSH) {
                par.group(words + ptr, np);
                for (uint32_t j = 0; j < np; ++j) call += ", " + par(words[ptr + j]);
            }
            call += ", " + q + "x[r], " + q + "y[r], " + q + "z[r])";
            body += "        float " + q + "x[R], " + q + "y[R], " + q + "z[R];\n        RM_FOR_R " + call + ";\n";
            pos.push_back(Pos{q + "x", q + "y", q + "z", op, words[ptr], stack.size()});
            ptr += np;
            continue;
        } else if (op == RM_OP_TRANSLATION_POP || op == RM_OP_ROTATION_POP || op == RM_OP_SCALE_POP) {
            if (pos.size() < 2 || pos.back().push_op + 1u != op || stack.size() != pos.back().depth + 1) {
                if (err) *err = "unbalanced transform group";
                return RM_E_STACK_UNBALANCED;
            }
            if (op == RM_OP_SCALE_POP) {
                if (!need(1)) { if (err) *err = "truncated scale pop"; return RM_E_STACK_UNBALANCED; }
                if (words[ptr] != pos.back().s_bits) { if (err) *err = "ScalePop factor differs from its push"; return RM_E_BAD_PARAM; }
                // the child's distance back in the parent's units: a new SSA value replaces the top of the stack
                const std::string top = var(stack.back()) + "[r]";
                body += "        float " + var(i) + "[R];\n        RM_FOR_R " + var(i) + "[r] = Arith<EX>::mul(" + top + ", " + par(words[ptr]) + ");\n";
                stack.back() = i;
                ptr += 1; est += kInsnScalePop;
            }
            pos.pop_back();
            continue;
        } else if (op == RM_OP_UNION || op == RM_OP_SUBTRACTION || op == RM_OP_INTERSECTION || op == RM_OP_SMOOTH_UNION) {
            if (stack.size() < 2 + pos.back().depth) { if (err) *err = "value stack underflow"; return RM_E_STACK_UNBALANCED; }
            const std::string b = var(stack.back()) + "[r]", b2 = wvar(stack.back()) + "[h]";  // b = pop, a = pop  (wgsl:243-244)
            stack.pop_back();
            const std::string a = var(stack.back()) + "[r]", a2 = wvar(stack.back()) + "[h]";
            stack.pop_back();
            if (op == RM_OP_UNION) { e = "fminf(" + a + ", " + b + ")"; e2 = "f2_min(" + a2 + ", " + b2 + ")"; est += kInsnMinMax; }
            else if (op == RM_OP_SUBTRACTION) { e = "fmaxf(" + a + ", -" + b + ")"; e2 = "f2_max(" + a2 + ", f2_neg(" + b2 + "))"; est += kInsnMinMax; }
            else if (op == RM_OP_INTERSECTION) { e = "fmaxf(" + a + ", " + b + ")"; e2 = "f2_max(" + a2 + ", " + b2 + ")"; est += kInsnMinMax; }
            else {
                if (!need(1)) { if (err) *err = "truncated smooth union"; return RM_E_STACK_UNBALANCED; }
                const float k = as_f32(words[ptr++]);
                if (!(k > 0.0f) || !std::isfinite(k)) { if (err) *err = "smooth-union k must be finite and > 0"; return RM_E_BAD_PARAM; }
                float inv_k, c4k;
                bool pow2;
                smooth_constants(k, &inv_k, &c4k, &pow2);
                if (literals && exact) {
                    e = "op_smooth<EX>(" + a + ", " + b + ", " + lit(as_u32(k)) + ", " + lit(as_u32(inv_k)) + ", " + lit(as_u32(c4k)) + ", " +
                        (pow2 ? "true" : "false") + ")";
                } else if (exact) {
                    // whether 1/k is exact decides multiply-or-divide: structural.  1/k is only read when it is.
                    if (sig && !literals) sig->push_back(0xF0000010u | (pow2 ? 1u : 0u));
                    const uint32_t kb[2] = {as_u32(k), as_u32(inv_k)};
                    par.group(kb, pow2 ? 2 : 1);
                    e = "op_smooth<EX>(" + a + ", " + b + ", " + par(kb[0]) + ", " + (pow2 ? par(kb[1]) : std::string("0.0f")) + ", 0.0f, " +
                        (pow2 ? "true" : "false") + ")";
                } else {
                    // fast mode reads (k*s, s) only (rm_program.h smooth_constants_fast)
                    float ks, sk;
                    smooth_constants_fast(k, &ks, &sk);
                    const uint32_t kb[2] = {as_u32(ks), as_u32(sk)};
                    par.group(kb, 2);
                    // k*s <= 1: the clamp at 0 is the FFMA's .SAT — structural (a k dragged across 4 is another kernel)
                    const bool sat = smooth_fast_sat_ok(ks);
                    if (sig && !literals) sig->push_back(0xF0000010u | (sat ? 2u : 0u));
                    const std::string ks_ = par(kb[0]), s_ = par(kb[1]);
                    e = std::string("op_smooth<EX, ") + (sat ? "true" : "false") + ">(" + a + ", " + b + ", 0.0f, " + ks_ + ", " + s_ + ", false)";
                    e2 = std::string("op_smooth2<") + (sat ? "true" : "false") + ">(" + a2 + ", " + b2 + ", " + ks_ + ", " + s_ + ")";
                }
                est += kInsnSmooth;
            }
            stack.push_back(i);
        } else {
            if (err) *err = "unknown opcode " + std::to_string(op);
            return RM_E_BAD_OPCODE;
        }
        if (stack.size() > RM_STACK_MAX) { if (err) *err = "value stack deeper than 32"; return RM_E_STACK_OVERFLOW; }
        body += "        float " + var(i) + "[R];\n        RM_FOR_R " + var(i) + "[r] = " + e + ";\n";
        if (packable) body2 += "        f2_t " + wvar(i) + "[R / 2];\n        RM_FOR_H " + wvar(i) + "[h] = " + e2 + ";\n";
    }
    // same acceptance as decode_program: at least one value left (the shader's final pop_value takes the top), all groups closed
    if (cmd_count != 0 && (stack.empty() || pos.size() != 1)) { if (err) *err = "value stack ends empty or a transform group is open"; return RM_E_STACK_UNBALANCED; }

    // rays per thread from the code-size estimate: the march loop's copy of map_scene has to stay in the instruction cache
    if (R == 0) R = env_int("RM_SPEC_R");
    if (R != 2 && R != 4 && R != 8) {
        if (est * 4 * kBytesPerInsn <= kLoopBudget4) R = 4;
        else if (est * 2 * kBytesPerInsn <= kLoopBudget2) R = 2;
        else {
            if (err) *err = "program too long to specialise usefully (" + std::to_string(est) + " instructions per ray): the interpreter is used";
            return RM_E_JIT;
        }
    }
    if ((size_t)par.n * sizeof(float) > kMaxParamBytes) {
        if (err) *err = "too many parameters for the kernel-argument block (" + std::to_string(par.n) + "): the interpreter is used";
        return RM_E_JIT;
    }
    // Deep programs (2 rays per thread) get ONE out-of-line copy of the scene code, called by the march loop and by the
    // tap loop, instead of an inlined copy in each: half the code (the instruction cache holds it with room to spare), no
    // spill at all, half the compile time, and specialisation keeps paying up to 64 leaves instead of 48 (measured on B200,
    // profiles/r2_s9_crossover.txt: 32 / 48 / 64 leaves x1.26 / 1.21 / 1.18 over the interpreter against x1.25 / 1.13 / 1.03
    // inlined).  Shallow programs (4 rays) stay inlined: the call costs more than it saves there (cfg2 0.708 vs 0.680 ms).
    // STRUCTURE tier of a deep program: an out-of-line function cannot read the kernel's arguments as constants (through a
    // reference they become generic loads: 38.3 ms against 37.0 inlined on cfg3), so the parameter block then lives in a
    // __constant__ variable of the generated module (g_rm_spec_P, filled by launch_spec before the launch whenever the
    // values change): LDCU / LDC from bank 3, one copy of the code — 35.8 ms (profiles/r2_s36_constbank.txt).
    bool noinline = R == 2;
    if (const char *ns = std::getenv("RM_SPEC_NOINLINE")) noinline = ns[0] == '1';
    bool constbank = noinline && !literals && par.n > 0;
    if (const char *cs = std::getenv("RM_SPEC_CONSTBANK")) constbank = constbank && cs[0] == '1';
    if (noinline && !literals && !constbank) noinline = std::getenv("RM_SPEC_NOINLINE") != nullptr;  // without the bank: inlined (measured)
    // pairs of rays through the packed FP32 instructions: fast mode, reference-set programs (RM_SPEC_PACKED=0: measurement switch)
    bool packed = packable && cmd_count != 0;
    if (const char *ps = std::getenv("RM_SPEC_PACKED")) packed = packed && ps[0] != '0';
    int minb = env_int("RM_SPEC_MINB");
    // measured on B200: profiles/r1_s14_NOTES.md, r2_s8_noinline_sweep.jsonl; in ray pairs (64-bit register pairs want more
    // registers) profiles/r2_s45_packed.txt: cfg3 8 / 9 / 10 / 11 / 12 blocks 28.04 / 27.27 / 27.10 / 26.92 / 26.92 ms, cfg2 8 / 9 /
    // 10 blocks 0.500 / 0.512 / 0.547 ms (10 spills inside the march loop)
    if (minb <= 0) {
        if (packed) minb = R == 2 ? (noinline ? (literals ? 11 : 10) : 12) : (R == 4 ? 8 : 7);  // structure tier: 28.47 ms at 10, 29.55 at 11
        else minb = R == 2 ? (noinline ? 10 : 12) : (R == 4 ? 10 : 7);
    }
    if (sig) { sig->push_back(0xF0000020u | (uint32_t)R); sig->push_back(0xF0000040u | (uint32_t)minb); sig->push_back(0xF0000080u | (noinline ? 1u : 0u) | (constbank ? 2u : 0u) | (packed ? 4u : 0u)); }
    if (r_out) *r_out = R;
    if (minb_out) *minb_out = minb;
    if (est_out) *est_out = est;
    if (constbank_out) *constbank_out = constbank;
    if (params) while (params->size() < 4 || params->size() % 4) params->push_back(0.0f);  // = sizeof(SpecParams): the launch copies all of it
    if (!src) return RM_OK;

    std::string s;
    s += "// generated by rm_jit.cu: map_scene of one " + std::to_string(cmd_count) + "-command stream as straight-line code\n";
    s += "#include \"rm_direct_body.cuh\"\n";
    s += "namespace rm {\n";
    {   // the parameter block's name in the scene code
        const std::string name = constbank ? "g_rm_spec_P" : "P";
        for (size_t at = body.find("@P@"); at != std::string::npos; at = body.find("@P@", at)) body.replace(at, 3, name);
        for (size_t at = body2.find("@P@"); at != std::string::npos; at = body2.find("@P@", at)) body2.replace(at, 3, name);
    }
    // the scene code of one evaluation: px / py / pz[R] in, out[R] out
    std::string scene;
    if (cmd_count == 0) scene = "        RM_FOR_R out[r] = max_dist;  // wgsl:189-191\n";
    else if (packed) {
        scene = "        f2_t PX[R / 2], PY[R / 2], PZ[R / 2];  // rays 2h and 2h+1 side by side: every FADD / FMUL / FFMA below issues once for both\n"
                "        RM_FOR_H { PX[h] = f2_pack(px[2 * h], px[2 * h + 1]); PY[h] = f2_pack(py[2 * h], py[2 * h + 1]); PZ[h] = f2_pack(pz[2 * h], pz[2 * h + 1]); }\n" +
                body2 + "        RM_FOR_H { out[2 * h] = f2_lo(" + wvar(stack.back()) + "[h]); out[2 * h + 1] = f2_hi(" + wvar(stack.back()) + "[h]); }\n";
    } else scene = body + "        RM_FOR_R out[r] = " + var(stack.back()) + "[r];\n";
    s += "struct alignas(16) SpecParams { float f[" + std::to_string(par.n ? ((par.n + 3u) & ~3u) : 4u) + "]; };  // the stream's non-zero parameters, in stream order\n";
    if (constbank) {
        s += "}  // namespace rm\n#define RM_SPEC_CONSTBANK 1\n";
        s += "extern \"C\" { __constant__ rm::SpecParams g_rm_spec_P; }  // written by launch_spec (a small kernel on the paint's stream) when the values change\n";
        s += "namespace rm {\n";
    }
    s += "template <bool EX, int R>\nstruct SpecMap {\n    float max_dist;\n";
    if (constbank) s += "    __device__ __forceinline__ SpecMap(float md, const SpecParams &) : max_dist(md) {}\n";
    else {
        s += "    const SpecParams &P;  // the kernel's __grid_constant__ argument\n";
        s += "    __device__ __forceinline__ SpecMap(float md, const SpecParams &p) : max_dist(md), P(p) {}\n";
    }
    if (noinline) {  // scalar arguments and a by-value result: nothing of the call goes through memory
        s += "    struct Out { float v[R]; };\n";
        s += "    __device__ __noinline__ Out eval(";
        for (int r = 0; r < R; ++r) s += std::string(r ? ", " : "") + "float x" + std::to_string(r) + ", float y" + std::to_string(r) + ", float z" + std::to_string(r);
        s += ") const {\n        float px[R], py[R], pz[R], out[R];\n";
        for (int r = 0; r < R; ++r) s += "        px[" + std::to_string(r) + "] = x" + std::to_string(r) + "; py[" + std::to_string(r) + "] = y" + std::to_string(r) + "; pz[" + std::to_string(r) + "] = z" + std::to_string(r) + ";\n";
        s += scene;
        s += "        Out o;\n        RM_FOR_R o.v[r] = out[r];\n        return o;\n    }\n";
        s += "    __device__ __forceinline__ void operator()(const float (&px)[R], const float (&py)[R], const float (&pz)[R], float (&out)[R]) {\n        const Out o = eval(";
        for (int r = 0; r < R; ++r) s += std::string(r ? ", " : "") + "px[" + std::to_string(r) + "], py[" + std::to_string(r) + "], pz[" + std::to_string(r) + "]";
        s += ");\n        RM_FOR_R out[r] = o.v[r];\n";
    } else {
    s += "    __device__ __forceinline__ void operator()(const float (&px)[R], const float (&py)[R], const float (&pz)[R], float (&out)[R]) {\n";
    s += scene;
    }
    s += "    }\n};\n}  // namespace rm\n";
    s += "extern \"C\" __global__ void __launch_bounds__(RM_DIRECT_THREADS, " + std::to_string(minb) + ")\n";
    s += "rm_spec(const __grid_constant__ rm::FrameParams fp, const __grid_constant__ rm::SpecParams P, uint8_t *__restrict__ dst, size_t pitch,\n        unsigned long long *__restrict__ counters, uint8_t *__restrict__ aov_cls, float *__restrict__ aov_t) {\n";
    s += "    rm::SpecMap<" + std::string(exact ? "true" : "false") + ", " + std::to_string(R) + "> map(fp.max_dist, P);\n";
    s += "    rm::direct_body<" + std::string(exact ? "true" : "false") + ", " + std::to_string(R) + ", " + (variant == SPEC_COUNT ? "true" : "false") + ", " +
         (variant == SPEC_AOV ? "true" : "false") + ">(fp, dst, pitch, counters, aov_cls, aov_t, map);\n}\n";
    *src = std::move(s);
    return RM_OK;
}

int spec_compile(const std::string &src, std::string *cubin, std::string *err) {
    const Nvrtc &n = nvrtc();
    if (!n.h) {
        if (err) *err = "NVRTC is not available (" + n.why + "): programs cannot be specialised, the interpreter is used";
        return RM_E_JIT;
    }
    const char *names[kNumHeaders], *texts[kNumHeaders];
    for (int i = 0; i < kNumHeaders; ++i) { names[i] = kHeaders[i].name; texts[i] = kHeaders[i].text; }
    nvrtcProgram prog = nullptr;
    int rc = n.CreateProgram(&prog, src.c_str(), "rm_spec.cu", kNumHeaders, texts, names);
    if (rc != 0) { if (err) *err = std::string("nvrtcCreateProgram: ") + n.GetErrorString(rc); return RM_E_JIT; }
    const char *opts[] = {"--gpu-architecture=sm_100a", "-std=c++17", "-lineinfo", "--fmad=true"};
    rc = n.CompileProgram(prog, (int)(sizeof(opts) / sizeof(opts[0])), opts);
    if (rc != 0) {
        size_t ls = 0;
        n.GetProgramLogSize(prog, &ls);
        std::string log(ls, '\0');
        if (ls) n.GetProgramLog(prog, &log[0]);
        if (err) *err = std::string("nvrtcCompileProgram: ") + n.GetErrorString(rc) + "\n" + log;
        n.DestroyProgram(&prog);
        return RM_E_JIT;
    }
    size_t cs = 0;
    n.GetCUBINSize(prog, &cs);
    cubin->assign(cs, '\0');
    rc = cs ? n.GetCUBIN(prog, &(*cubin)[0]) : 1;
    n.DestroyProgram(&prog);
    if (rc != 0) { if (err) *err = "nvrtcGetCUBIN failed"; return RM_E_JIT; }
    return RM_OK;
}

int spec_get(uint32_t cmd_count, const uint32_t *words, uint32_t n_words, bool exact, int variant, bool literals, SpecHandle *out,
             std::vector<float> *params, bool *compiled_now, std::string *err, bool lookup_only) {
    if (compiled_now) *compiled_now = false;
    Key key{{}, exact, variant};
    std::string src, gen_err;
    SpecKernel meta;
    // generation is cheap (string building); it yields the structure signature = the cache key, and the parameter block
    int rc = spec_generate_source(cmd_count, words, n_words, exact, variant, 0, &src, &meta.rays_per_thread, &meta.min_blocks,
                                  &meta.est_insn_per_ray, &gen_err, params, &key.sig, literals, &meta.constbank);
    if (rc != RM_OK) {
        if (err) *err = gen_err;
        return rc;
    }
    // The cache mutex is NOT held while compiling (a paint that only looks a kernel up must never wait for somebody else's
    // compile): a miss inserts a placeholder, compiles outside the lock and fills it in; requests for the same key wait
    // for it (blocking callers) or report a miss (lookup_only).
    std::unique_lock<std::mutex> lock(g_mu);
    auto it = g_cache.find(key);
    if (it != g_cache.end() && !it->second.ready) {
        if (lookup_only) return RM_E_JIT;
        g_cv.wait(lock, [&] { auto i2 = g_cache.find(key); return i2 == g_cache.end() || i2->second.ready; });
        it = g_cache.find(key);
    }
    if (it == g_cache.end()) {
        if (lookup_only) return RM_E_JIT;
        while (g_cache.size() >= g_cache_cap) {
            if (!evict_one()) {
                if (err) *err = "specialisation cache is full (" + std::to_string(g_cache_cap) + " programs, all in use): the interpreter is used";
                return RM_E_JIT;
            }
        }
        if (compiled_now) *compiled_now = true;
        g_cache.emplace(key, Entry{});  // placeholder: ready = false
        lock.unlock();
        Entry e;
        const auto t0 = std::chrono::steady_clock::now();
        std::string cubin;
        auto k = std::make_shared<SpecKernel>();
        k->exact = exact; k->variant = variant; k->literals = literals; k->constbank = meta.constbank;
        k->rays_per_thread = meta.rays_per_thread; k->min_blocks = meta.min_blocks; k->est_insn_per_ray = meta.est_insn_per_ray;
        k->n_params = params ? (uint32_t)params->size() : 0;
        e.rc = spec_compile(src, &cubin, &e.err);
        if (e.rc == RM_OK) {
            cudaError_t ce = cudaLibraryLoadData(&k->lib, cubin.data(), nullptr, nullptr, 0, nullptr, nullptr, 0);
            if (ce == cudaSuccess) ce = cudaLibraryGetKernel(&k->fn, k->lib, "rm_spec");
            if (ce != cudaSuccess) {
                e.rc = RM_E_JIT;
                e.err = std::string("loading the specialised kernel: ") + cudaGetErrorString(ce);
                cudaGetLastError();
            }
            k->cubin_bytes = cubin.size();
        }
        k->compile_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        if (e.rc == RM_OK) e.k = std::move(k);
        e.ready = true;
        lock.lock();
        it = g_cache.find(key);
        if (it == g_cache.end()) it = g_cache.emplace(key, Entry{}).first;  // (evicted meanwhile: cannot happen to a placeholder, kept for safety)
        it->second = std::move(e);  // failures are cached too: no retry storm per frame
        g_cv.notify_all();
    }
    Entry &e = it->second;
    e.last_use = ++g_tick;
    if (e.rc != RM_OK) {
        if (err) *err = e.err;
        return e.rc;
    }
    *out = e.k;
    return RM_OK;
}

namespace {
// the refill of a module's __constant__ parameter block: the values travel as a kernel argument (no host buffer has to
// outlive the call, no host synchronisation), the stores go through the symbol's global address
struct ParamBlock { float f[kMaxParamBytes / sizeof(float)]; };
__global__ void k_set_spec_params(float *__restrict__ dst, const __grid_constant__ ParamBlock p, uint32_t n) {
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) dst[i] = p.f[i];
}
}  // namespace

cudaError_t launch_spec(const SpecKernel &k, const FrameParams &fp, const float *params, uint8_t *dst, size_t pitch,
                        unsigned long long *counters, uint8_t *aov_cls, float *aov_t, cudaStream_t st) {
    if (fp.nrows == 0 || fp.W == 0) return cudaSuccess;
    const uint32_t tile_h = direct_tile_rows(k.rays_per_thread);
    FrameParams p = fp;
    static const float kNoParams[4] = {0.0f, 0.0f, 0.0f, 0.0f};
    const float *pp = params ? params : kNoParams;  // the launch copies sizeof(SpecParams) bytes from here
    int dev = 0;
    cudaGetDevice(&dev);
    k.dev_mask.fetch_or(1ull << (dev & 63));
    void *args[] = {&p, (void *)pp, &dst, &pitch, &counters, &aov_cls, &aov_t};
    if (!k.constbank) return cudaLaunchKernel((const void *)k.fn, direct_grid(fp, tile_h), dim3(RM_DIRECT_THREADS), args, 0, st);

    std::lock_guard<std::mutex> lock(k.cb_mu);
    SpecKernel::ConstBank &b = k.cb[dev];
    cudaError_t e = cudaSuccess;
    if (!b.sym) {
        void *sym = nullptr;
        size_t bytes = 0;
        e = cudaLibraryGetGlobal(&sym, &bytes, k.lib, "g_rm_spec_P");
        if (e == cudaSuccess && bytes < (size_t)k.n_params * sizeof(float)) e = cudaErrorInvalidSymbol;
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&b.last_launch, cudaEventDisableTiming);
        if (e != cudaSuccess) return e;
        b.sym = (float *)sym;
    }
    // launches of this kernel on this device run one after the other, whatever their streams
    if (b.launched && b.last_stream != st && (e = cudaStreamWaitEvent(st, b.last_launch, 0)) != cudaSuccess) return e;
    if (b.written.size() != k.n_params || memcmp(b.written.data(), pp, (size_t)k.n_params * sizeof(float)) != 0) {
        ParamBlock blk;
        memcpy(blk.f, pp, (size_t)k.n_params * sizeof(float));
        k_set_spec_params<<<1, 256, 0, st>>>(b.sym, blk, k.n_params);
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
        b.written.assign(pp, pp + k.n_params);
    }
    e = cudaLaunchKernel((const void *)k.fn, direct_grid(fp, tile_h), dim3(RM_DIRECT_THREADS), args, 0, st);
    if (e != cudaSuccess) return e;
    b.launched = true;
    b.last_stream = st;
    return cudaEventRecord(b.last_launch, st);
}

}  // namespace rm
