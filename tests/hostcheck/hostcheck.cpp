// hostcheck.cpp — TEST INFRASTRUCTURE ONLY.  Compiles the product's host-side program decoder (rm_program.cpp) and its
// DEVICE-side interpreters (rm_device.cuh map_eval1, rm_wavefront_common.cuh map_evalR) for the host through
// shim/cuda_runtime.h, so that tests/test_hostcheck_cpu.py can compare what the kernels would compute per map_scene
// evaluation with the oracle without a GPU (exact mode: bit for bit).  Built by the test into tests/hostcheck/_build;
// never shipped, never loaded by ray-marching_b200.
#include "cuda_runtime.h"  // the shim

#include "../../include/rm_b200.h"
#include "../../ray-marching_b200/csrc/rm_program.cpp"
#include "../../ray-marching_b200/csrc/rm_wavefront_common.cuh"

using namespace rm;

namespace {

template <bool EX, bool EXT>
void run1(const Program &p, float max_dist, const float *pts, uint64_t n, float *out) {
    const float4 *prog = reinterpret_cast<const float4 *>(p.insns.data());
    for (uint64_t i = 0; i < n; ++i)
        out[i] = map_eval1<EX, EXT>(prog, (uint32_t)p.insns.size(), max_dist, pts[3 * i], pts[3 * i + 1], pts[3 * i + 2]);
}

template <bool EX, int R, bool EXT>
void runR(const Program &p, float max_dist, const float *pts, uint64_t n, float *out) {
    const float4 *prog = reinterpret_cast<const float4 *>(p.insns.data());
    float spill[RM_STACK_MAX - 1][R];
    for (uint64_t i = 0; i < n; i += R) {
        float px[R], py[R], pz[R], s[R];
        for (int r = 0; r < R; ++r) {
            const uint64_t j = i + r < n ? i + r : n - 1;
            px[r] = pts[3 * j]; py[r] = pts[3 * j + 1]; pz[r] = pts[3 * j + 2];
        }
        map_evalR<EX, R, false, EXT>(prog, (uint32_t)p.insns.size(), max_dist, px, py, pz, s, spill);
        for (int r = 0; r < R && i + r < n; ++r) out[i + r] = s[r];
    }
}

}  // namespace

extern "C" {

// decoded instruction stream (48-byte RmInsn records) of a command stream; returns the RM_* code of rm_prepare
int hc_decode(uint32_t cmd_count, const uint32_t *words, uint32_t n_words, uint32_t max_words, void *insns, uint32_t cap,
              uint32_t *n_insn, int *extended, uint32_t *flops_per_eval, int collapse_chains) {
    Program p;
    const int rc = decode_program(cmd_count, words, n_words, max_words, &p, collapse_chains != 0);
    if (rc != RM_OK) return rc;
    if (n_insn) *n_insn = (uint32_t)p.insns.size();
    if (extended) *extended = p.extended ? 1 : 0;
    if (flops_per_eval) *flops_per_eval = p.flops_per_eval();
    if (insns) memcpy(insns, p.insns.data(), sizeof(RmInsn) * std::min<size_t>(cap, p.insns.size()));
    return RM_OK;
}

// map_scene at n points through the interpreter the kernels run: rays_per_thread 1 = map_eval1 (verification kernels),
// 4 / 8 = map_evalR (production kernel).  force_ext runs a reference-set program through the EXT instantiation too.
// collapse_chains: the experimental fast-mode lowering of rm_program.h (one similarity per primitive).
int hc_map(int exact, int rays_per_thread, int force_ext, uint32_t cmd_count, const uint32_t *words, uint32_t n_words, float max_dist,
           const float *pts, uint64_t n, float *out, int collapse_chains) {
    Program p;
    const int rc = decode_program(cmd_count, words, n_words, 3072, &p, collapse_chains != 0);
    if (rc != RM_OK) return rc;
    const bool ext = p.extended || force_ext;
#define HC_RUN(EXV, EXTV)                                                            \
    do {                                                                             \
        if (rays_per_thread == 1) run1<EXV, EXTV>(p, max_dist, pts, n, out);         \
        else if (rays_per_thread == 4) runR<EXV, 4, EXTV>(p, max_dist, pts, n, out); \
        else if (rays_per_thread == 8) runR<EXV, 8, EXTV>(p, max_dist, pts, n, out); \
        else return RM_E_INVALID_ARG;                                                \
    } while (0)
    if (exact) { if (ext) HC_RUN(true, true); else HC_RUN(true, false); }
    else { if (ext) HC_RUN(false, true); else HC_RUN(false, false); }
#undef HC_RUN
    return RM_OK;
}

}  // extern "C"
