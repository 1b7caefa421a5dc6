// rm_jit.h — program specialisation: map_scene of ONE command stream emitted as straight-line CUDA, compiled at run
// time (NVRTC, sm_100a) around the production kernel body (rm_direct_body.cuh), cached process-wide.
//
// Why (measured on B200, profiles/r1_s14_NOTES.md): the decoded-program interpreter issues 168 warp instructions per
// fused instruction for 8 rays, 23 of them fetch/dispatch/copies, and holds the instruction's constants in registers.
// Straight-line code has none of that — constants are immediates, there is no dispatch — and needs fewer registers, so
// more warps are resident.  What bounds it instead is the instruction cache: a 63-command scene is 9 KB of code per ray,
// so the rays per thread drop from 8 to 2 to keep the march loop inside the 32-KB L1.5 instruction cache
// (8 rays: 80 KB loop, 61 ms; 4 rays: 40 KB, 42.4 ms; 2 rays: 18 KB, 34.5 ms; interpreter 44.0 ms).
// Arithmetic per ray is operation for operation the interpreter's: exact mode stays bit-identical to the oracle.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <memory>
#include <string>
#include <vector>

#include <map>
#include <mutex>
#include <vector>

#include "rm_device.cuh"

namespace rm {

// what the kernel does besides painting: nothing, the work counters (rm_set_counters), or the per-sample AOVs (rm_paint_aov)
enum { SPEC_PLAIN = 0, SPEC_COUNT = 1, SPEC_AOV = 2 };

struct SpecKernel {
    cudaKernel_t fn = nullptr;
    cudaLibrary_t lib = nullptr;
    int rays_per_thread = 0;   // 2, 4 or 8
    int min_blocks = 0;        // __launch_bounds__ second argument
    bool exact = false;
    int variant = 0;           // SPEC_PLAIN / SPEC_COUNT / SPEC_AOV
    double compile_ms = 0.0;   // NVRTC compile + module load
    size_t cubin_bytes = 0;
    uint32_t est_insn_per_ray = 0;
    uint32_t n_params = 0;     // floats of the SpecParams kernel argument
    bool literals = false;     // literal tier: every parameter an immediate, one kernel per word stream
    // Structure tier of a deep program: the scene code is out of line and reads its parameters from the module's
    // __constant__ copy g_rm_spec_P (one per device).  launch_spec refills it — a small kernel on the paint's stream —
    // whenever the values differ from the ones last written on that device, and orders the launches of this kernel on a
    // device one after the other across streams, so that no launch still reads the block while it is rewritten (two
    // contexts of one device may paint the same structure with different values).
    bool constbank = false;
    struct ConstBank {
        float *sym = nullptr;            // device address of g_rm_spec_P on this device
        std::vector<float> written;      // the values last written (n_params floats)
        cudaEvent_t last_launch = nullptr;
        cudaStream_t last_stream = nullptr;
        bool launched = false;
    };
    mutable std::mutex cb_mu;
    mutable std::map<int, ConstBank> cb;  // by device
    mutable std::atomic<uint64_t> dev_mask{0};  // devices that have launched it (synchronised before the cubin is unloaded)
    SpecKernel() = default;
    SpecKernel(const SpecKernel &) = delete;
    SpecKernel &operator=(const SpecKernel &) = delete;
    ~SpecKernel();  // unloads the cubin
};
// Contexts hold the kernels of their current program by shared_ptr: the cache may evict an entry (LRU) while a context
// still paints with it; the cubin is unloaded when the last holder lets go.
using SpecHandle = std::shared_ptr<const SpecKernel>;

// CUDA source of the specialised kernel for this (already validated) command stream.  rays_per_thread 0 = choose from
// the code-size estimate.  Returns RM_OK, or RM_E_JIT when the program is too long to specialise usefully.
// Two tiers.  STRUCTURE tier (literals = false): the code depends on the stream's structure only — opcode sequence, which
// parameters are exactly zero, and in exact mode which smooth-union k are powers of two; `sig` receives that signature
// (the cache key) and `params` the values the kernel reads from its SpecParams argument (the non-zero parameters in
// stream order, grouped per command on 16 / 8-byte boundaries; smooth unions contribute their derived constants, ScalePush
// its 1/s).  One compile serves every edit of a value.  LITERAL tier (literals = true): every parameter an immediate by
// its bits, `sig` is the whole word stream, `params` comes back as zeros: ~5 % fewer instructions on the deep scene
// (no LDCU loads), worth compiling for words that stay put.
int spec_generate_source(uint32_t cmd_count, const uint32_t *words, uint32_t n_words, bool exact, int variant,
                         int rays_per_thread, std::string *src, int *r_out, int *minb_out, uint32_t *est_out, std::string *err,
                         std::vector<float> *params = nullptr, std::vector<uint32_t> *sig = nullptr, bool literals = false,
                         bool *constbank_out = nullptr);

// NVRTC-compiles `src` for sm_100a to a cubin (no GPU needed).  RM_E_JIT + message when libnvrtc cannot be loaded or
// the compile fails.
int spec_compile(const std::string &src, std::string *cubin, std::string *err);

// Cached: generate + (on a signature miss) compile + load for (program structure, mode, variant); `params` receives the
// parameter block of THESE words.  `compiled_now` reports whether this call had to compile.  lookup_only: never compile —
// RM_E_JIT on a miss (the background policy uses it to adopt a cached kernel without starting a thread).
int spec_get(uint32_t cmd_count, const uint32_t *words, uint32_t n_words, bool exact, int variant, bool literals, SpecHandle *out,
             std::vector<float> *params, bool *compiled_now, std::string *err, bool lookup_only = false);

// LRU bound of the process-wide cache (default 512 structures) and its counters.
void spec_cache_set_capacity(size_t n);
void spec_cache_stats(size_t *entries, size_t *capacity, uint64_t *evictions);

// params: k.n_params floats (the block spec_get / spec_generate_source produced for the words being painted)
cudaError_t launch_spec(const SpecKernel &k, const FrameParams &fp, const float *params, uint8_t *dst, size_t pitch,
                        unsigned long long *counters, uint8_t *aov_cls, float *aov_t, cudaStream_t st);

}  // namespace rm
