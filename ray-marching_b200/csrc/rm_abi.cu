// rm_abi.cu — the extern "C" entry points of include/rm_b200.h.
// Host-side state of one device: the reference's RayMarchingResources (renderer.rs:43-49) minus
// all wgpu plumbing — limits, command buffer (decoded), uniforms — plus a framebuffer.
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <cmath>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <thread>

#include "../../include/rm_b200.h"
#include "rm_cull.h"
#include "rm_direct_body.cuh"
#include "rm_frame.h"
#include "rm_jit.h"
#include "rm_kernels.h"
#include "rm_program.h"
#include "rm_wavefront.h"

using namespace rm;

namespace {
constexpr uint32_t kMaxWordsCap = 3072;  // decoded program must fit in shared memory (48 B/insn)
constexpr int kEventSlots = 16;
constexpr size_t kFlushBytes = 256ull << 20;
}  // namespace

struct rm_ctx {
    int device = 0;
    int sm_count = 0;
    uint32_t max_words = RM_DEFAULT_MAX_WORDS;
    cudaStream_t stream = nullptr;
    cudaEvent_t events[kEventSlots] = {};
    cudaEvent_t upload_done = nullptr;
    // recorded behind every paint on the stream it was launched on: the kernels stage d_prog block by block for as long
    // as they run, so the next program upload (ctx stream) has to wait for it
    cudaEvent_t paint_done = nullptr;
    bool paint_pending = false;
    int mode = RM_MODE_EXACT;
    int kernel = RM_KERNEL_DIRECT;
    bool prepared = false;
    bool counters_on = false;
    bool early_out = false;  // rm_set_early_out: opt-in, never on by default
    // rm_set_cull: opt-in per-tile culling of the program (rm_cull.cu); like the early-out it paints through the interpreter
    bool cull = false;
    unsigned long long *d_masks = nullptr;  // one per block tile of the band being painted (ctx-wide scratch)
    size_t masks_cap = 0;
    float cull_t_end = 0.0f;                // per program + uniforms: beyond this ray parameter no ray of the frame is marching
    uint64_t cull_paints = 0;
    // program specialisation (rm_jit.cu): paint the DIRECT kernel through run-time compiled straight-line code
    int spec_policy = 0;          // rm_set_specialize: 1 = every program, compiled at its first paint; 2 = compiled in the background
    // policy 2: one worker thread per context compiles while the interpreter keeps painting
    std::thread spec_worker;
    std::atomic<int> spec_bg_state{0};  // 0 idle, 1 compiling, 2 result ready
    uint64_t spec_bg_generation = 0;    // prog_uploads of the words being compiled: a result for older words is dropped
    int spec_bg_ex = 0, spec_bg_variant = 0, spec_bg_rc = 0;
    bool spec_bg_compiled = false, spec_bg_literals = false;
    std::string spec_bg_err;
    bool spec_requested = false;  // rm_specialize: the current program and every later one of the same STRUCTURE
    SpecHandle spec_cur[2][3];    // [exact][SPEC_PLAIN / SPEC_COUNT / SPEC_AOV] for the current words
    bool spec_failed[2][3] = {};
    // the kernels are compiled per program structure (rm_jit.h): the values travel as a kernel argument
    std::vector<float> spec_params;      // parameter block of the current words (the same for both modes and all variants)
    // LITERAL tier (rm_jit.h): the current words with every parameter an immediate — adopted once the same words have
    // been painted spec_lit_after times (compiled in the background, never a stall) or at once by rm_specialize
    SpecHandle spec_lit[2][3];
    bool spec_lit_failed[2][3] = {};
    uint32_t spec_lit_after = 8;         // rm_set_spec_literal_after: 0 = structure tier only
    uint64_t paints_same_words = 0;      // plain paints since the words last changed
    std::vector<uint32_t> spec_sig[2];   // structure signature the explicit request (rm_specialize) was made for, per mode
    std::string spec_status;      // why the last specialisation failed
    std::atomic<uint64_t> spec_compiles{0};  // also counted by the background worker, when its compile finishes
    uint64_t spec_launches = 0;
    float limits_f[2] = {0.01f, 100.0f};  // renderer.rs:133-137
    uint32_t max_iter = 100;
    float uniforms[36] = {};
    // The command stream decoded twice (rm_program.h): `prog` keeps the oracle's sequential operation order (RM_MODE_EXACT),
    // `prog_fast` is the algebraically equal form RM_MODE_FAST paints (smooth unions as min(a,b) - e*e; the transform chains
    // of extended programs composed into one similarity per primitive).  Both live in d_prog: exact at 0, fast at prog_cap.
    Program prog, prog_fast;
    RmInsn *d_prog = nullptr;
    RmInsn *h_prog = nullptr;  // pinned staging
    size_t prog_cap = 0;       // instructions per form
    uint8_t *d_fb = nullptr;
    size_t fb_cap = 0, fb_bytes = 0;
    float *d_codes = nullptr;
    size_t codes_cap = 0;
    void *d_flush = nullptr;
    unsigned long long *d_counters = nullptr;
    WavefrontState wf;
    cudaStream_t last_paint_stream = nullptr;
    // rm_paint_poses_host: device frame ring + copy stream so that frame p's D2H overlaps frame p+1's render
    static constexpr int kRing = 3;
    cudaStream_t copy_stream = nullptr;
    // rm_paint_poses: odd frames run on a second stream so that one frame's tail overlaps the next frame's start
    cudaStream_t pose_stream = nullptr;
    cudaEvent_t pose_fork = nullptr, pose_join = nullptr;
    uint8_t *d_ring = nullptr;
    size_t ring_cap = 0;
    cudaEvent_t ring_rendered[kRing] = {}, ring_copied[kRing] = {};
    uint64_t launches = 0;
    // last uploaded command stream: rm_prepare re-uses the decoded program on the device when the words are unchanged
    std::vector<uint32_t> last_words;
    uint32_t last_cmd_count = 0;
    uint64_t prog_uploads = 0;
    // shared frames (rm_shared_frame_*): allocations this ctx owns, and peer allocations it has opened over CUDA IPC
    std::vector<void *> shared_owned, shared_opened;
    std::string err;
};

namespace {

int fail(rm_ctx *c, int code, const char *what, cudaError_t e = cudaSuccess) {
    if (c) {
        c->err = what;
        if (e != cudaSuccess) {
            c->err += ": ";
            c->err += cudaGetErrorString(e);
        }
    }
    return code;
}
#define RM_CUDA(ctx, call)                                             \
    do {                                                               \
        cudaError_t e__ = (call);                                      \
        if (e__ != cudaSuccess) return fail((ctx), RM_E_CUDA, #call, e__); \
    } while (0)

int ensure(rm_ctx *c, void **p, size_t *cap, size_t need) {
    if (need <= *cap) return RM_OK;
    if (*p) cudaFree(*p);
    *p = nullptr;
    *cap = 0;
    RM_CUDA(c, cudaMalloc(p, need));
    *cap = need;
    return RM_OK;
}

// the decoded form the current arithmetic mode paints, on the host and on the device
const Program &cur_prog(const rm_ctx *c) { return c->mode == RM_MODE_EXACT ? c->prog : c->prog_fast; }
const RmInsn *cur_dprog(const rm_ctx *c) { return c->mode == RM_MODE_EXACT ? c->d_prog : c->d_prog + c->prog_cap; }

FrameParams make_params(const rm_ctx *c, const float *u, uint32_t W, uint32_t H, uint32_t row0, uint32_t nrows) {
    return make_frame_params(u, c->limits_f[0], c->limits_f[1], c->max_iter, (uint32_t)cur_prog(c).insns.size(), W, H, row0, nrows);
}

int check_band(rm_ctx *c, uint32_t W, uint32_t H, uint32_t row0, uint32_t nrows) {
    if (!c) return RM_E_INVALID_ARG;
    if (W == 0 || H == 0 || W > 65536 || H > 65536) return fail(c, RM_E_INVALID_ARG, "bad frame size");
    if ((uint64_t)row0 + nrows > H) return fail(c, RM_E_INVALID_ARG, "band outside frame");
    if (!c->prepared) return fail(c, RM_E_NOT_PREPARED, "rm_prepare has not been called");
    return RM_OK;
}

}  // namespace

extern "C" {

const char *rm_version(void) { return "rm_b200 0.1 (sm_100a)"; }

const char *rm_strerror(int code) {
    switch (code) {
        case RM_OK: return "ok";
        case RM_E_INVALID_ARG: return "invalid argument";
        case RM_E_PROGRAM_TOO_LARGE: return "command buffer larger than capacity";
        case RM_E_BAD_OPCODE: return "unknown command opcode";
        case RM_E_STACK_OVERFLOW: return "value stack deeper than 32 (or transform groups nested deeper than 8)";
        case RM_E_STACK_UNBALANCED: return "value stack underflow, truncated command stream or unbalanced transform group";
        case RM_E_CUDA: return "CUDA error";
        case RM_E_NOT_PREPARED: return "no program/uniforms prepared";
        case RM_E_BAD_PARAM: return "bad command parameter";
        case RM_E_JIT: return "program specialisation failed (NVRTC unavailable, compile error or program too long)";
        case RM_E_UNSUPPORTED: return "the selected kernel does not run this program";
        default: return "unknown error";
    }
}

const char *rm_last_error(rm_ctx *ctx) { return ctx ? ctx->err.c_str() : "null ctx"; }

int rm_create_ex(int cuda_device, uint32_t max_words, rm_ctx **out) {
    if (!out || max_words == 0 || max_words > kMaxWordsCap) return RM_E_INVALID_ARG;
    *out = nullptr;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || cuda_device < 0 || cuda_device >= n) return RM_E_CUDA;
    if (cudaSetDevice(cuda_device) != cudaSuccess) return RM_E_CUDA;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, cuda_device) != cudaSuccess) return RM_E_CUDA;
    if (prop.major != 10) return RM_E_CUDA;  // kernels are built for sm_100a only; no fallback
    rm_ctx *c = new (std::nothrow) rm_ctx();
    if (!c) return RM_E_INVALID_ARG;
    c->device = cuda_device;
    c->sm_count = prop.multiProcessorCount;
    c->max_words = max_words;
    bool ok = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) == cudaSuccess;
    for (int i = 0; ok && i < kEventSlots; ++i) ok = cudaEventCreate(&c->events[i]) == cudaSuccess;
    ok = ok && cudaEventCreateWithFlags(&c->upload_done, cudaEventDisableTiming) == cudaSuccess;
    ok = ok && cudaEventCreateWithFlags(&c->paint_done, cudaEventDisableTiming) == cudaSuccess;
    c->prog_cap = max_words;
    ok = ok && cudaMalloc(&c->d_prog, 2 * c->prog_cap * sizeof(RmInsn)) == cudaSuccess;
    ok = ok && cudaMallocHost(&c->h_prog, 2 * c->prog_cap * sizeof(RmInsn)) == cudaSuccess;
    ok = ok && cudaMalloc(&c->d_counters, 8 * sizeof(unsigned long long)) == cudaSuccess;  // 4 work counters + executed flops (culled paints)
    ok = ok && cudaMemset(c->d_counters, 0, 8 * sizeof(unsigned long long)) == cudaSuccess;
    ok = ok && wavefront_init(&c->wf, cuda_device) == cudaSuccess;
    if (!ok) {
        rm_destroy(c);
        return RM_E_CUDA;
    }
    *out = c;
    return RM_OK;
}

int rm_create(int cuda_device, rm_ctx **out) { return rm_create_ex(cuda_device, RM_DEFAULT_MAX_WORDS, out); }

void rm_destroy(rm_ctx *c) {
    if (!c) return;
    if (c->spec_worker.joinable()) c->spec_worker.join();
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    wavefront_destroy(&c->wf);
    for (auto &e : c->events) if (e) cudaEventDestroy(e);
    if (c->upload_done) cudaEventDestroy(c->upload_done);
    if (c->paint_done) cudaEventDestroy(c->paint_done);
    if (c->d_prog) cudaFree(c->d_prog);
    if (c->h_prog) cudaFreeHost(c->h_prog);
    for (void *p : c->shared_opened) cudaIpcCloseMemHandle(p);
    for (void *p : c->shared_owned) cudaFree(p);
    if (c->d_fb) cudaFree(c->d_fb);
    if (c->copy_stream) { cudaStreamSynchronize(c->copy_stream); cudaStreamDestroy(c->copy_stream); }
    if (c->pose_stream) { cudaStreamSynchronize(c->pose_stream); cudaStreamDestroy(c->pose_stream); }
    if (c->pose_fork) cudaEventDestroy(c->pose_fork);
    if (c->pose_join) cudaEventDestroy(c->pose_join);
    if (c->d_ring) cudaFree(c->d_ring);
    for (auto &e : c->ring_rendered) if (e) cudaEventDestroy(e);
    for (auto &e : c->ring_copied) if (e) cudaEventDestroy(e);
    if (c->d_codes) cudaFree(c->d_codes);
    if (c->d_masks) cudaFree(c->d_masks);
    if (c->d_flush) cudaFree(c->d_flush);
    if (c->d_counters) cudaFree(c->d_counters);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

int rm_set_limits(rm_ctx *c, const void *limits12) {
    if (!c || !limits12) return RM_E_INVALID_ARG;
    float f[2];
    uint32_t it;
    std::memcpy(f, limits12, 8);
    std::memcpy(&it, (const char *)limits12 + 8, 4);
    c->limits_f[0] = f[0]; c->limits_f[1] = f[1]; c->max_iter = it;
    return RM_OK;
}

int rm_set_mode(rm_ctx *c, int mode) {
    if (!c || (mode != RM_MODE_EXACT && mode != RM_MODE_FAST)) return RM_E_INVALID_ARG;
    c->mode = mode;  // both decoded forms of the program are on the device (rm_prepare): nothing to redo
    return RM_OK;
}
int rm_get_mode(rm_ctx *c) { return c ? c->mode : RM_E_INVALID_ARG; }

int rm_validate_program(uint32_t cmd_count, const uint32_t *words, uint32_t n_words, uint32_t max_words) {
    return decode_program(cmd_count, words, n_words, max_words, nullptr);
}

int rm_prepare_uniforms(rm_ctx *c, const void *uniforms144) {
    if (!c || !uniforms144) return RM_E_INVALID_ARG;
    std::memcpy(c->uniforms, uniforms144, RM_UNIFORMS_BYTES);
    return RM_OK;
}

int rm_prepare(rm_ctx *c, const void *uniforms144, uint32_t cmd_count, const uint32_t *words, uint32_t n_words) {
    if (!c || !uniforms144) return RM_E_INVALID_ARG;
    if (n_words && !words) return fail(c, RM_E_INVALID_ARG, "null words");
    // Same command stream as the one already on the device (the reference re-flattens and re-uploads the tree every
    // frame, renderer.rs:224-239; an editor frame usually changes only the camera): only the uniforms change.
    if (c->prepared && cmd_count == c->last_cmd_count && n_words == c->last_words.size() &&
        (n_words == 0 || std::memcmp(words, c->last_words.data(), (size_t)n_words * sizeof(uint32_t)) == 0)) {
        std::memcpy(c->uniforms, uniforms144, RM_UNIFORMS_BYTES);
        return RM_OK;
    }
    Program p, pf;
    int rc = decode_program(cmd_count, words, n_words, c->max_words, &p);
    if (rc == RM_OK) rc = decode_program(cmd_count, words, n_words, c->max_words, &pf, /*fast_form=*/true, /*collapse_chains=*/true);
    if (rc != RM_OK) return fail(c, rc, rm_strerror(rc));
    // an extended program (N4) re-establishes the evaluation point whenever the chain of transforms in force changes, so
    // its instruction stream can be longer than its word stream; the device buffer holds max_words instructions
    if (p.insns.size() > c->prog_cap || pf.insns.size() > c->prog_cap) return fail(c, RM_E_PROGRAM_TOO_LARGE, "decoded program is longer than the instruction capacity (rm_create_ex max_words)");
    RM_CUDA(c, cudaSetDevice(c->device));
    // the staging buffer may still be in flight from the previous prepare
    RM_CUDA(c, cudaEventSynchronize(c->upload_done));
    c->prog = std::move(p);
    c->prog_fast = std::move(pf);
    const size_t n = c->prog.insns.size(), nf = c->prog_fast.insns.size();
    // paints still running on a caller-supplied stream read d_prog until they finish: the upload goes behind them
    if (c->paint_pending) {
        RM_CUDA(c, cudaStreamWaitEvent(c->stream, c->paint_done, 0));
        c->paint_pending = false;
    }
    if (n) {
        std::memcpy(c->h_prog, c->prog.insns.data(), n * sizeof(RmInsn));
        RM_CUDA(c, cudaMemcpyAsync(c->d_prog, c->h_prog, n * sizeof(RmInsn), cudaMemcpyHostToDevice, c->stream));
    }
    if (nf) {
        std::memcpy(c->h_prog + c->prog_cap, c->prog_fast.insns.data(), nf * sizeof(RmInsn));
        RM_CUDA(c, cudaMemcpyAsync(c->d_prog + c->prog_cap, c->h_prog + c->prog_cap, nf * sizeof(RmInsn), cudaMemcpyHostToDevice, c->stream));
    }
    RM_CUDA(c, cudaEventRecord(c->upload_done, c->stream));
    std::memcpy(c->uniforms, uniforms144, RM_UNIFORMS_BYTES);
    c->last_words.assign(words, words + n_words);
    c->last_cmd_count = cmd_count;
    // A specialised kernel belongs to one program STRUCTURE (opcode sequence; rm_jit.h): new words of the same structure —
    // an editor dragging a radius — find their kernel in the cache at the next paint and only the parameter block changes.
    // An explicit rm_specialize (no policy) holds for that structure only: other words go back to the interpreter.
    if (c->spec_requested && !c->spec_policy) {
        const int ex = c->mode == RM_MODE_EXACT ? 1 : 0;
        std::vector<uint32_t> sig;
        if (c->spec_sig[ex].empty() ||
            spec_generate_source(cmd_count, words, n_words, ex != 0, SPEC_PLAIN, 0, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, &sig) != RM_OK ||
            sig != c->spec_sig[ex])
            c->spec_requested = false;
    }
    for (int a = 0; a < 2; ++a) for (int b = 0; b < 3; ++b) {
        c->spec_cur[a][b].reset(); c->spec_failed[a][b] = false;
        c->spec_lit[a][b].reset(); c->spec_lit_failed[a][b] = false;
    }
    c->paints_same_words = 0;
    if (!c->spec_requested) { c->spec_sig[0].clear(); c->spec_sig[1].clear(); }
    ++c->prog_uploads;
    c->prepared = true;
    return RM_OK;
}

uint64_t rm_program_uploads(rm_ctx *c) { return c ? c->prog_uploads : 0; }

}  // extern "C"

// paint / debug / measurement entry points live in rm_abi_paint.inc to keep files short
#include "rm_abi_paint.inc"
