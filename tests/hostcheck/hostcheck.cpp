// hostcheck.cpp — TEST INFRASTRUCTURE ONLY.  Compiles the product's host-side program decoder (rm_program.cpp) and its
// DEVICE-side interpreters (rm_device.cuh map_eval1, rm_wavefront_common.cuh map_evalR) for the host through
// shim/cuda_runtime.h, so that tests/test_hostcheck_cpu.py can compare what the kernels would compute per map_scene
// evaluation with the oracle without a GPU (exact mode: bit for bit).  Built by the test into tests/hostcheck/_build;
// never shipped, never loaded by ray-marching_b200.
#include "hc_grid.h"  // the shim (cuda_runtime.h) + the lane-per-thread grid runner

#include "../../include/rm_b200.h"
#include "../../ray-marching_b200/csrc/rm_program.cpp"
#include "../../ray-marching_b200/csrc/rm_wavefront_common.cuh"
#include "../../ray-marching_b200/csrc/rm_direct_body.cuh"
#include "../../ray-marching_b200/csrc/rm_frame.h"

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

// The production kernel body (rm_direct_body.cuh) around the interpreter, as k_direct / k_direct_aov of rm_direct.cu run
// it: one block = 128 threads = an 8-wide pixel tile, grid = one block per tile.
template <bool EX, int R, bool COUNT, bool AOV, bool EXT>
void paint_t(const FrameParams &fp, const Program &p, uint8_t *dst, size_t pitch, unsigned long long *counters, uint8_t *cls, float *t) {
    const float4 *prog = reinterpret_cast<const float4 *>(p.insns.data());
    const uint32_t tile_h = direct_tile_rows(R);
    hc_run_grid(direct_grid(fp, tile_h), RM_DIRECT_THREADS, [&] {
        InterpMap<EX, R, false, EXT> map;
        map.prog = prog; map.n_insn = fp.n_insn; map.max_dist = fp.max_dist;
        direct_body<EX, R, COUNT, AOV>(fp, dst, pitch, counters, cls, t, map);
    });
}

}  // namespace

extern "C" {

// A band of a frame through the production kernel body + interpreter, on the host.  rays_per_thread 0 = the library's
// rule (4 up to 4 decoded instructions, else 8).  counters4 != NULL: the COUNT variant; aov_cls / aov_t != NULL: the AOV one.
int hc_paint(int exact, int rays_per_thread, uint32_t cmd_count, const uint32_t *words, uint32_t n_words, const void *uniforms144,
             const void *limits12, uint32_t W, uint32_t H, uint32_t row0, uint32_t nrows, uint32_t strip_rows, uint32_t strip_pitch,
             int dst_frame_rows, uint8_t *dst, size_t pitch, uint64_t *counters4, uint8_t *aov_cls, float *aov_t) {
    Program p;
    // exact < 0: fast mode with the collapsed lowering of extended programs (the fast form rm_prepare decodes)
    const int rc = decode_program(cmd_count, words, n_words, 3072, &p, /*fast_form=*/exact <= 0, /*collapse_chains=*/exact < 0);
    if (exact < 0) exact = 0;
    if (rc != RM_OK) return rc;
    float u[36], lim[2];
    uint32_t max_iter;
    memcpy(u, uniforms144, 144);
    memcpy(lim, limits12, 8);
    memcpy(&max_iter, (const char *)limits12 + 8, 4);
    FrameParams fp = make_frame_params(u, lim[0], lim[1], max_iter, (uint32_t)p.insns.size(), W, H, row0, nrows);
    fp.strip_rows = strip_rows; fp.strip_pitch = strip_pitch; fp.dst_frame_rows = dst_frame_rows ? 1u : 0u;
    if (nrows == 0 || W == 0) return RM_OK;
    const int R = rays_per_thread ? rays_per_thread : (fp.n_insn <= 4 ? 4 : 8);
    if (R != 4 && R != 8) return RM_E_INVALID_ARG;
    unsigned long long *cnt = reinterpret_cast<unsigned long long *>(counters4);
    const bool aov = aov_cls && aov_t;
#define HC_PAINT3(EXV, RV, EXTV)                                                                        \
    do {                                                                                                \
        if (aov) paint_t<EXV, RV, false, true, EXTV>(fp, p, dst, pitch, nullptr, aov_cls, aov_t);       \
        else if (cnt) paint_t<EXV, RV, true, false, EXTV>(fp, p, dst, pitch, cnt, nullptr, nullptr);    \
        else paint_t<EXV, RV, false, false, EXTV>(fp, p, dst, pitch, nullptr, nullptr, nullptr);        \
    } while (0)
#define HC_PAINT2(EXV, RV) do { if (p.extended) HC_PAINT3(EXV, RV, true); else HC_PAINT3(EXV, RV, false); } while (0)
    if (exact) { if (R == 8) HC_PAINT2(true, 8); else HC_PAINT2(true, 4); }
    else { if (R == 8) HC_PAINT2(false, 8); else HC_PAINT2(false, 4); }
#undef HC_PAINT2
#undef HC_PAINT3
    return RM_OK;
}

// decoded instruction stream (48-byte RmInsn records) of a command stream; returns the RM_* code of rm_prepare
int hc_decode(uint32_t cmd_count, const uint32_t *words, uint32_t n_words, uint32_t max_words, void *insns, uint32_t cap,
              uint32_t *n_insn, int *extended, uint32_t *flops_per_eval, int collapse_chains) {
    Program p;
    const int rc = decode_program(cmd_count, words, n_words, max_words, &p, /*fast_form=*/collapse_chains != 0, collapse_chains != 0);
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
    const int rc = decode_program(cmd_count, words, n_words, 3072, &p, /*fast_form=*/!exact, collapse_chains != 0);
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
