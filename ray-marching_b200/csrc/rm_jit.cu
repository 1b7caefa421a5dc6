// rm_jit.cu — see rm_jit.h.  Emits map_scene of one command stream as straight-line CUDA, compiles it with NVRTC
// (loaded with dlopen: the library has no link-time dependency on it) for sm_100a and loads the cubin through the
// runtime's library API.  Everything device-side comes from the same headers the ahead-of-time kernels use; they are
// embedded in this object as strings (rm_jit_headers.inc, written by the Makefile) and handed to NVRTC by name.
#include "rm_jit.h"

#include <dlfcn.h>

#include <chrono>
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
constexpr uint32_t kInsnSphere = 8, kInsnBox = 17, kInsnMinMax = 1, kInsnSmooth = 6;
// N4 commands, counted from their device functions (not yet measured): plane 3 FFMA + FMUL-free, translation 3 FADD,
// rotation 3 FMUL + 6 FFMA, scale 3 FMUL (push) + 1 FMUL (pop)
constexpr uint32_t kInsnPlane = 3, kInsnTranslate = 3, kInsnRotate = 9, kInsnScalePush = 3, kInsnScalePop = 1;
constexpr uint32_t kBytesPerInsn = 16;
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
};

const Nvrtc &nvrtc() {
    static const Nvrtc n = [] {
        Nvrtc r;
        std::vector<std::string> names;
        if (const char *p = std::getenv("RM_NVRTC_PATH")) names.push_back(p);  // exact file; nothing else is tried
        else {
            names = {"libnvrtc.so.12", "libnvrtc.so", "/usr/local/cuda/lib64/libnvrtc.so.12", "/usr/local/cuda/lib64/libnvrtc.so"};
            if (const char *home = std::getenv("CUDA_HOME")) names.push_back(std::string(home) + "/lib64/libnvrtc.so");
        }
        for (const auto &nm : names) {
            r.h = dlopen(nm.c_str(), RTLD_NOW | RTLD_LOCAL);
            if (r.h) break;
            const char *de = dlerror();
            r.why += nm + ": " + (de ? de : "not loadable") + "; ";
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

// ---- process-wide cache ------------------------------------------------------------------------
struct Key {
    std::vector<uint32_t> words;
    uint32_t cmd_count;
    bool exact;
    int variant;
    bool operator<(const Key &o) const {
        if (cmd_count != o.cmd_count) return cmd_count < o.cmd_count;
        if (exact != o.exact) return exact < o.exact;
        if (variant != o.variant) return variant < o.variant;
        return words < o.words;
    }
};
struct Entry {
    int rc = RM_OK;
    std::string err;
    SpecKernel k;
};
constexpr size_t kMaxCacheEntries = 512;
std::mutex g_mu;
std::map<Key, std::unique_ptr<Entry>> g_cache;

}  // namespace

int spec_generate_source(uint32_t cmd_count, const uint32_t *words, uint32_t n_words, bool exact, int variant, int R,
                         std::string *src, int *r_out, int *minb_out, uint32_t *est_out, std::string *err) {
    // walk the postfix stream (builder.rs:26-62) with a stack of SSA value numbers: command i defines v<i>
    std::string body;
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
        std::string e;
        if (op == RM_OP_SPHERE) {
            if (!need(4)) { if (err) *err = "truncated sphere"; return RM_E_STACK_UNBALANCED; }
            e = "sd_sphere<EX>(" + at() + ", " + lit(words[ptr]) + ", " + lit(words[ptr + 1]) + ", " + lit(words[ptr + 2]) +
                ", " + lit(words[ptr + 3]) + ")";
            ptr += 4; est += kInsnSphere;
            stack.push_back(i);
        } else if (op == RM_OP_BOX) {
            if (!need(6)) { if (err) *err = "truncated box"; return RM_E_STACK_UNBALANCED; }
            e = "sd_box<EX>(" + at();
            for (int j = 0; j < 6; ++j) e += ", " + lit(words[ptr + j]);
            e += ")";
            ptr += 6; est += kInsnBox;
            stack.push_back(i);
        } else if (op == RM_OP_PLANE) {
            if (!need(4)) { if (err) *err = "truncated plane"; return RM_E_STACK_UNBALANCED; }
            e = "sd_plane<EX>(" + at();
            for (int j = 0; j < 4; ++j) e += ", " + lit(words[ptr + j]);
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
                if (!(sc > 0.0f) || sc > 3.0e38f) { if (err) *err = "scale must be finite and > 0"; return RM_E_BAD_PARAM; }
                call = "xf_scale<EX>(" + at() + ", " + lit(words[ptr]) + ", " + lit(as_u32(1.0f / sc));
                est += kInsnScalePush;
            }
            if (op != RM_OP_SCALE_PUSH) for (uint32_t j = 0; j < np; ++j) call += ", " + lit(words[ptr + j]);
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
                body += "        float " + var(i) + "[R];\n        RM_FOR_R " + var(i) + "[r] = Arith<EX>::mul(" + top + ", " + lit(words[ptr]) + ");\n";
                stack.back() = i;
                ptr += 1; est += kInsnScalePop;
            }
            pos.pop_back();
            continue;
        } else if (op == RM_OP_UNION || op == RM_OP_SUBTRACTION || op == RM_OP_INTERSECTION || op == RM_OP_SMOOTH_UNION) {
            if (stack.size() < 2 + pos.back().depth) { if (err) *err = "value stack underflow"; return RM_E_STACK_UNBALANCED; }
            const std::string b = var(stack.back()) + "[r]";  // b = pop, a = pop  (wgsl:243-244)
            stack.pop_back();
            const std::string a = var(stack.back()) + "[r]";
            stack.pop_back();
            if (op == RM_OP_UNION) { e = "fminf(" + a + ", " + b + ")"; est += kInsnMinMax; }
            else if (op == RM_OP_SUBTRACTION) { e = "fmaxf(" + a + ", -" + b + ")"; est += kInsnMinMax; }
            else if (op == RM_OP_INTERSECTION) { e = "fmaxf(" + a + ", " + b + ")"; est += kInsnMinMax; }
            else {
                if (!need(1)) { if (err) *err = "truncated smooth union"; return RM_E_STACK_UNBALANCED; }
                const float k = as_f32(words[ptr++]);
                if (!(k > 0.0f) || k != k || k > 3.0e38f) { if (err) *err = "smooth-union k must be finite and > 0"; return RM_E_BAD_PARAM; }
                float inv_k, c4k;
                bool pow2;
                smooth_constants(k, &inv_k, &c4k, &pow2);
                e = "op_smooth<EX>(" + a + ", " + b + ", " + lit(as_u32(k)) + ", " + lit(as_u32(inv_k)) + ", " + lit(as_u32(c4k)) + ", " +
                    (pow2 ? "true" : "false") + ")";
                est += kInsnSmooth;
            }
            stack.push_back(i);
        } else {
            if (err) *err = "unknown opcode " + std::to_string(op);
            return RM_E_BAD_OPCODE;
        }
        if (stack.size() > RM_STACK_MAX) { if (err) *err = "value stack deeper than 32"; return RM_E_STACK_OVERFLOW; }
        body += "        float " + var(i) + "[R];\n        RM_FOR_R " + var(i) + "[r] = " + e + ";\n";
    }
    if (cmd_count != 0 && (stack.size() != 1 || pos.size() != 1)) { if (err) *err = "value stack does not end with one value"; return RM_E_STACK_UNBALANCED; }

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
    int minb = env_int("RM_SPEC_MINB");
    if (minb <= 0) minb = R == 2 ? 12 : (R == 4 ? 10 : 7);  // measured on B200, profiles/r1_s14_NOTES.md
    if (r_out) *r_out = R;
    if (minb_out) *minb_out = minb;
    if (est_out) *est_out = est;
    if (!src) return RM_OK;

    std::string s;
    s += "// generated by rm_jit.cu: map_scene of one " + std::to_string(cmd_count) + "-command stream as straight-line code\n";
    s += "#include \"rm_direct_body.cuh\"\n";
    s += "namespace rm {\n";
    s += "template <bool EX, int R>\nstruct SpecMap {\n    float max_dist;\n";
    s += "    __device__ __forceinline__ void operator()(const float (&px)[R], const float (&py)[R], const float (&pz)[R], float (&out)[R]) {\n";
    if (cmd_count == 0) s += "        RM_FOR_R out[r] = max_dist;  // wgsl:189-191\n";
    else s += body + "        RM_FOR_R out[r] = " + var(stack.back()) + "[r];\n";
    s += "    }\n};\n}  // namespace rm\n";
    s += "extern \"C\" __global__ void __launch_bounds__(RM_DIRECT_THREADS, " + std::to_string(minb) + ")\n";
    s += "rm_spec(const __grid_constant__ rm::FrameParams fp, uint8_t *__restrict__ dst, size_t pitch, unsigned long long *__restrict__ counters,\n        uint8_t *__restrict__ aov_cls, float *__restrict__ aov_t) {\n";
    s += "    rm::SpecMap<" + std::string(exact ? "true" : "false") + ", " + std::to_string(R) + "> map;\n    map.max_dist = fp.max_dist;\n";
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

int spec_get(uint32_t cmd_count, const uint32_t *words, uint32_t n_words, bool exact, int variant, const SpecKernel **out,
             bool *compiled_now, std::string *err) {
    Key key{std::vector<uint32_t>(words, words + n_words), cmd_count, exact, variant};
    std::lock_guard<std::mutex> lock(g_mu);  // one compile at a time; the other members of a group then hit the cache
    if (compiled_now) *compiled_now = false;
    auto it = g_cache.find(key);
    if (it == g_cache.end() && g_cache.size() >= kMaxCacheEntries) {
        // kernels are never unloaded (contexts hold pointers into the cache), so the cache is bounded instead: a host that
        // streams thousands of distinct programs with the policy on falls back to the interpreter for the new ones
        if (err) *err = "specialisation cache is full (" + std::to_string(kMaxCacheEntries) + " programs): the interpreter is used";
        return RM_E_JIT;
    }
    if (it == g_cache.end()) {
        if (compiled_now) *compiled_now = true;
        auto e = std::make_unique<Entry>();
        const auto t0 = std::chrono::steady_clock::now();
        std::string src, cubin;
        e->k.exact = exact; e->k.variant = variant;
        e->rc = spec_generate_source(cmd_count, words, n_words, exact, variant, 0, &src, &e->k.rays_per_thread, &e->k.min_blocks,
                                     &e->k.est_insn_per_ray, &e->err);
        if (e->rc == RM_OK) e->rc = spec_compile(src, &cubin, &e->err);
        if (e->rc == RM_OK) {
            cudaError_t ce = cudaLibraryLoadData(&e->k.lib, cubin.data(), nullptr, nullptr, 0, nullptr, nullptr, 0);
            if (ce == cudaSuccess) ce = cudaLibraryGetKernel(&e->k.fn, e->k.lib, "rm_spec");
            if (ce != cudaSuccess) {
                e->rc = RM_E_JIT;
                e->err = std::string("loading the specialised kernel: ") + cudaGetErrorString(ce);
                cudaGetLastError();
            }
            e->k.cubin_bytes = cubin.size();
        }
        e->k.compile_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        it = g_cache.emplace(std::move(key), std::move(e)).first;  // failures are cached too: no retry storm per frame
    }
    const Entry &e = *it->second;
    if (e.rc != RM_OK) {
        if (err) *err = e.err;
        return e.rc;
    }
    *out = &e.k;
    return RM_OK;
}

cudaError_t launch_spec(const SpecKernel &k, const FrameParams &fp, uint8_t *dst, size_t pitch, unsigned long long *counters,
                        uint8_t *aov_cls, float *aov_t, cudaStream_t st) {
    if (fp.nrows == 0 || fp.W == 0) return cudaSuccess;
    const uint32_t tile_h = direct_tile_rows(k.rays_per_thread);
    const unsigned grid = ((fp.W + RM_DIRECT_TILE_W - 1) / RM_DIRECT_TILE_W) * ((fp.nrows + tile_h - 1) / tile_h);
    FrameParams p = fp;
    void *args[] = {&p, &dst, &pitch, &counters, &aov_cls, &aov_t};
    return cudaLaunchKernel((const void *)k.fn, dim3(grid), dim3(RM_DIRECT_THREADS), args, 0, st);
}

}  // namespace rm
