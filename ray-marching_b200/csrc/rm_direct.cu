// rm_direct.cu — the production kernel of the render pass, ahead-of-time flavour: no ray queues, no wavefront
// bookkeeping at all.  The kernel body (ray generation, lock-step march of R sample rays per thread, normals, in-order
// resolve, 128-bit stores) is rm_direct_body.cuh; here map_scene is the R-wide decoded-program INTERPRETER
// (rm_wavefront_common.cuh map_evalR): one instruction fetch and one flag dispatch feed R evaluations.
// Works for any program with no set-up cost.  rm_jit.cu compiles the same body around straight-line code for ONE program.
//
// Measured on B200 (profiles/r1_s4_NOTES.md): for 1-3 instruction programs the wavefront kernels spend more on ray
// queues, refill and tile accounting than on map_scene, and even the one-thread-per-sample verification kernel beats
// them (S0 0.169 vs 0.377 ms).  That kernel in turn pays the interpreter's fetch/dispatch once per ray; this one
// amortises it over 4 or 8 rays.  It was written for shallow programs first and measured faster than the wavefront
// kernels on the deep scene too.
#include "rm_wavefront.h"
#include "rm_wavefront_common.cuh"
#include "rm_direct_body.cuh"

#include <cstdlib>

namespace rm {

namespace {

constexpr int DIRECT_THREADS = RM_DIRECT_THREADS;
constexpr int TILE_W_D = RM_DIRECT_TILE_W;
// R = sample rays per thread (4 or 8 here; the run-time specialised kernels of rm_jit.cu also use 2); the body is
// rm_direct_body.cuh.
// Resident blocks per SM asked of ptxas (measured on B200, profiles/r1_s12_*):
//   8-ray variant, cfg3: 4 blocks (123 registers, no spill) 45.99 ms, 5: 45.41, 6: 45.19 -> after the ray-generation change
//   6: 44.76, 7 (72 registers, 196 B spill): 44.02, 8 (64 registers): 45.13.  One more warp per scheduler beats the spill up to 7.
//   4-ray variant, cfg2: no hint (70 registers) 1.203 ms, 8 blocks (64 registers, 44 B spill) 1.184, 9: 1.189, 10: 1.232;
//   an explicit hint of 1 let ptxas take 77 registers and cost 6 %.
#ifndef RM_DIRECT_MINB8
#define RM_DIRECT_MINB8 7
#endif
#ifndef RM_DIRECT_MINB4
#define RM_DIRECT_MINB4 8
#endif
// Extended programs (N4: plane, space transforms) also carry the evaluation point q of every ray (3 R registers) through
// the interpreter loop.  Measured on B200 at 1080p (tools/n4_timing.py, profiles/r1_s17_*): 8-ray variant on scene S3 with
// 4 / 5 / 6 / 7 resident blocks 21.47 / 21.44 / 21.54 / 20.73 ms; 4-ray variant on a 3-instruction program with
// 6 / 8 / 10 blocks 2.877 / 2.905 / 3.069 ms.
#ifndef RM_DIRECT_MINB8_EXT
#define RM_DIRECT_MINB8_EXT 7
#endif
#ifndef RM_DIRECT_MINB4_EXT
#define RM_DIRECT_MINB4_EXT 6
#endif
constexpr int direct_min_blocks(int R, bool EXT) {
    return EXT ? (R == 8 ? RM_DIRECT_MINB8_EXT : RM_DIRECT_MINB4_EXT) : (R == 8 ? RM_DIRECT_MINB8 : RM_DIRECT_MINB4);
}

// algorithmic flops of one decoded instruction, for the executed-flop counter of culled paints
__device__ __forceinline__ uint32_t insn_flops(uint32_t fl) {
    uint32_t f = 0;
    if (!(fl & FLAG_POP)) f += (fl & FLAG_BOX) ? 22u : 10u;
    if (fl & FLAG_SMOOTH) f += 10u;
    else if (fl & FLAG_SUB) f += 2u;
    else if (fl & (FLAG_UNION | FLAG_INTER)) f += 1u;
    return f;
}

// CULL (opt-in, rm_cull.cu): the block stages only the instructions its tile's mask keeps — a shorter program, the same bits.
template <bool EX, int R, bool COUNT, bool EO, bool EXT, bool CULL = false>
__global__ void __launch_bounds__(DIRECT_THREADS, direct_min_blocks(R, EXT))
k_direct(const __grid_constant__ FrameParams fp, const RmInsn *__restrict__ prog_g, uint8_t *__restrict__ dst, size_t pitch,
         unsigned long long *__restrict__ counters, const unsigned long long *__restrict__ masks) {
    extern __shared__ float4 s_prog[];
    InterpMap<EX, R, EO, EXT> map;
    uint32_t eval_flops = 0;
    if constexpr (CULL) {
        __shared__ uint32_t s_flops;
        const unsigned long long mask = masks[blockIdx.y * gridDim.x + blockIdx.x];  // row-major over the tiles, as k_cull_masks wrote them
        if (threadIdx.x == 0) s_flops = 0;
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < fp.n_insn; i += blockDim.x) {
            if ((mask >> i) & 1ull) {
                const uint32_t pos = (uint32_t)__popcll(mask & ((1ull << i) - 1ull));
                const float4 *src = reinterpret_cast<const float4 *>(prog_g) + 3 * i;
                s_prog[3 * pos] = src[0]; s_prog[3 * pos + 1] = src[1]; s_prog[3 * pos + 2] = src[2];
                if (COUNT) atomicAdd(&s_flops, insn_flops(__float_as_uint(src[2].w)));
            }
        }
        __syncthreads();
        map.n_insn = (uint32_t)__popcll(mask);
        if (COUNT) eval_flops = s_flops;
    } else {
        const uint32_t nq = fp.n_insn * 3;
        for (uint32_t i = threadIdx.x; i < nq; i += blockDim.x) s_prog[i] = reinterpret_cast<const float4 *>(prog_g)[i];
        __syncthreads();
        map.n_insn = fp.n_insn;
    }
    map.prog = s_prog; map.max_dist = fp.max_dist;
    direct_body<EX, R, COUNT, false>(fp, dst, pitch, counters, nullptr, nullptr, map, eval_flops);
}

// the same kernel with the per-sample AOVs written out (verification: rm_paint_aov)
template <bool EX, int R, bool EXT>
__global__ void __launch_bounds__(DIRECT_THREADS, direct_min_blocks(R, EXT))
k_direct_aov(const __grid_constant__ FrameParams fp, const RmInsn *__restrict__ prog_g, uint8_t *__restrict__ dst, size_t pitch,
             uint8_t *__restrict__ aov_cls, float *__restrict__ aov_t) {
    extern __shared__ float4 s_prog[];
    const uint32_t nq = fp.n_insn * 3;
    for (uint32_t i = threadIdx.x; i < nq; i += blockDim.x) s_prog[i] = reinterpret_cast<const float4 *>(prog_g)[i];
    __syncthreads();
    InterpMap<EX, R, false, EXT> map;
    map.prog = s_prog; map.n_insn = fp.n_insn; map.max_dist = fp.max_dist;
    direct_body<EX, R, false, true>(fp, dst, pitch, nullptr, aov_cls, aov_t, map);
}

template <bool EX, int R, bool EXT>
cudaError_t launch_direct_aov_t(const FrameParams &fp, const RmInsn *prog, uint8_t *dst, size_t pitch, uint8_t *cls, float *t,
                                cudaStream_t st) {
    auto kern = k_direct_aov<EX, R, EXT>;
    const size_t smem = (size_t)fp.n_insn * sizeof(RmInsn);
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    const uint32_t tile_h = direct_tile_rows(R);
    kern<<<direct_grid(fp, tile_h), DIRECT_THREADS, smem, st>>>(fp, prog, dst, pitch, cls, t);
    return cudaGetLastError();
}

template <bool EX, int R, bool COUNT, bool EO, bool EXT, bool CULL = false>
cudaError_t launch_direct_t(const FrameParams &fp, const RmInsn *prog, uint8_t *dst, size_t pitch,
                            unsigned long long *counters, cudaStream_t st, const unsigned long long *masks = nullptr) {
    auto kern = k_direct<EX, R, COUNT, EO, EXT, CULL>;
    constexpr uint32_t tile_h = direct_tile_rows(R);
    const size_t smem = (size_t)fp.n_insn * sizeof(RmInsn);
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    kern<<<direct_grid(fp, tile_h), DIRECT_THREADS, smem, st>>>(fp, prog, dst, pitch, counters, masks);
    return cudaGetLastError();
}

}  // namespace

static int direct_rays_per_thread(const FrameParams &fp);

cudaError_t launch_direct_aov(bool exact, bool extended, const FrameParams &fp, const RmInsn *prog_dev, uint8_t *dst, size_t pitch,
                              uint8_t *cls, float *t, cudaStream_t st) {
    if (fp.nrows == 0 || fp.W == 0) return cudaSuccess;
    const int kR = direct_rays_per_thread(fp);
#define RM_AOV_DISPATCH(EXV, RV) \
    (extended ? launch_direct_aov_t<EXV, RV, true>(fp, prog_dev, dst, pitch, cls, t, st) : launch_direct_aov_t<EXV, RV, false>(fp, prog_dev, dst, pitch, cls, t, st))
    if (exact) return kR == 8 ? RM_AOV_DISPATCH(true, 8) : RM_AOV_DISPATCH(true, 4);
    return kR == 8 ? RM_AOV_DISPATCH(false, 8) : RM_AOV_DISPATCH(false, 4);
#undef RM_AOV_DISPATCH
}

cudaError_t launch_direct(bool exact, bool early_out, bool extended, const FrameParams &fp, const RmInsn *prog_dev, uint8_t *dst,
                          size_t pitch, unsigned long long *counters, cudaStream_t st) {
    if (fp.nrows == 0 || fp.W == 0) return cudaSuccess;
    const int kR = direct_rays_per_thread(fp);
#define RM_DIRECT_DISPATCH2(EXV, RV, EOV, EXTV)                                                      \
    (counters ? launch_direct_t<EXV, RV, true, EOV, EXTV>(fp, prog_dev, dst, pitch, counters, st)     \
              : launch_direct_t<EXV, RV, false, EOV, EXTV>(fp, prog_dev, dst, pitch, counters, st))
    // extended programs (N4) run the EXT interpreter; the early-out is written for the reference instruction set only
#define RM_DIRECT_DISPATCH(EXV, RV) \
    (extended ? RM_DIRECT_DISPATCH2(EXV, RV, false, true) : (early_out ? RM_DIRECT_DISPATCH2(EXV, RV, true, false) : RM_DIRECT_DISPATCH2(EXV, RV, false, false)))
    if (exact) return kR == 8 ? RM_DIRECT_DISPATCH(true, 8) : RM_DIRECT_DISPATCH(true, 4);
    return kR == 8 ? RM_DIRECT_DISPATCH(false, 8) : RM_DIRECT_DISPATCH(false, 4);
#undef RM_DIRECT_DISPATCH2
#undef RM_DIRECT_DISPATCH
}

int direct_rays_for(const FrameParams &fp) { return direct_rays_per_thread(fp); }

// The culled paint (rm_cull.cu): reference-set programs of at most 64 instructions, no early-out.  masks: one per block tile.
cudaError_t launch_direct_culled(bool exact, const FrameParams &fp, const RmInsn *prog_dev, uint8_t *dst, size_t pitch,
                                 unsigned long long *counters, const unsigned long long *masks, cudaStream_t st) {
    if (fp.nrows == 0 || fp.W == 0) return cudaSuccess;
    const int kR = direct_rays_per_thread(fp);
#define RM_CULL_DISPATCH(EXV, RV)                                                                                 \
    (counters ? launch_direct_t<EXV, RV, true, false, false, true>(fp, prog_dev, dst, pitch, counters, st, masks)  \
              : launch_direct_t<EXV, RV, false, false, false, true>(fp, prog_dev, dst, pitch, counters, st, masks))
    if (exact) return kR == 8 ? RM_CULL_DISPATCH(true, 8) : RM_CULL_DISPATCH(true, 4);
    return kR == 8 ? RM_CULL_DISPATCH(false, 8) : RM_CULL_DISPATCH(false, 4);
#undef RM_CULL_DISPATCH
}

// Rays per thread: 4 for shallow programs (more threads, less lock-step waste), 8 for deep ones (fetch/dispatch
// amortised over more rays).  Measured on B200 on truncated S2 chains at 1080p (tools/crossover.py, ms per frame,
// R=4 / R=8): 1 instruction 0.72/0.85, 3: 1.38/1.47, 4: 1.80/1.82, 5: 2.10/2.09, 8: 3.21/3.14, 32: 12.39/11.87.
// RM_DIRECT_R overrides.
static int direct_rays_per_thread(const FrameParams &fp) {
    static const int kForce = [] {
        const char *s = std::getenv("RM_DIRECT_R");
        const int v = s ? std::atoi(s) : 0;
        return (v == 4 || v == 8) ? v : 0;
    }();
    return kForce ? kForce : (fp.n_insn <= 4 ? 4 : 8);
}

}  // namespace rm
