// rm_direct.cu — the production kernel of the render pass: no ray queues, no wavefront bookkeeping at all.
// (Written for shallow programs first; measured faster than the wavefront kernels on the deep scene too.)
//
// Measured on B200 (profiles/r1_s4_NOTES.md): for 1-3 instruction programs the wavefront kernels spend more on ray
// queues, refill and tile accounting than on map_scene, and even the one-thread-per-sample verification kernel beats
// them (S0 0.169 vs 0.377 ms).  That kernel in turn pays the interpreter's fetch/dispatch once per ray.  Here a thread
// owns the 4 sample rays of one pixel that share the x offset i (samples 4i..4i+3, wgsl:44-45) and marches them in
// lock-step through the R-wide interpreter (wgsl:87-115): dispatch is amortised over 4 rays, the rays are a fraction of
// a pixel apart so they leave the loop within a step or two of each other, and the warp (8 neighbouring pixels) is
// coherent.  Rays that hit get their normal one ray at a time — the 4 tetrahedron taps (wgsl:138-142) fill the same 4
// slots — and the 16 samples of a pixel are summed in the shader's order by a shuffle chain over the pixel's 4 threads
// (wgsl:68-75).  Per-ray arithmetic is the same operation for operation as in the other kernels: bit-identical output.
#include "rm_wavefront.h"
#include "rm_wavefront_common.cuh"

#include <cstdlib>

namespace rm {

namespace {

constexpr int DIRECT_THREADS = 128;
// R = sample rays per thread (4 or 8); a pixel's 16 samples are held by TPP = 16/R neighbouring lanes, so a warp covers
// 32/TPP pixels = ROWS rows of 8, and the 4-warp block an 8 x (4*ROWS) pixel tile.
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
template <bool EX, int R, bool COUNT, bool EO>
__global__ void __launch_bounds__(DIRECT_THREADS, (R == 8 ? RM_DIRECT_MINB8 : RM_DIRECT_MINB4))
k_direct(const __grid_constant__ FrameParams fp, const RmInsn *__restrict__ prog_g, uint8_t *__restrict__ dst, size_t pitch,
         unsigned long long *__restrict__ counters) {
    static_assert(R == 4 || R == 8, "4 or 8 rays per thread");
    constexpr int TPP = 16 / R;          // threads per pixel
    constexpr int ROWS = (32 / TPP) / 8; // tile rows per warp
    using A = Arith<EX>;
    extern __shared__ float4 s_prog[];
    const uint32_t nq = fp.n_insn * 3;
    for (uint32_t i = threadIdx.x; i < nq; i += blockDim.x) s_prog[i] = reinterpret_cast<const float4 *>(prog_g)[i];
    __syncthreads();

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t tiles_x = (fp.W + TILE_W - 1) / TILE_W;
    const uint32_t tx = blockIdx.x % tiles_x, ty = blockIdx.x / tiles_x;
    const uint32_t pw = (uint32_t)lane / TPP;  // pixel of the warp, row-major over ROWS rows of 8
    const uint32_t px = tx * TILE_W + (pw & 7u), ry = (ty * 4u + (uint32_t)warp) * ROWS + (pw >> 3);
    const uint32_t grp = (uint32_t)lane % TPP;  // this thread holds samples grp*R .. grp*R+R-1 (sample = 4*i + j)
    const bool inside = px < fp.W && ry < fp.nrows;

    float ro[3] = {0.f, 0.f, 0.f};
    float rdx[R], rdy[R], rdz[R], dist[R], code[R];
    uint32_t live = 0, hit = 0;  // bit r: ray r still marching / ray r hit the scene
    uint32_t c_step = 0, c_hit = 0, c_floor = 0, c_sample = 0;
    RM_FOR_R { rdx[r] = rdy[r] = rdz[r] = 0.f; dist[r] = 0.f; code[r] = RM_CODE_SKY; }
    if (inside) {
        const uint32_t fy = frame_row(fp, ry);
        RM_FOR_R {
            float o[3], d[3], c[3], t;
            gen_ray<EX>(fp, px, fy, grp * R + (uint32_t)r, o, d);  // wgsl:39-62
            ro[0] = o[0]; ro[1] = o[1]; ro[2] = o[2];
            rdx[r] = d[0]; rdy[r] = d[1]; rdz[r] = d[2];
            const int cls = shade_miss(o, d, c, t);  // wgsl:117-130: what the ray gets if it never hits
            code[r] = cls == 0 ? RM_CODE_SKY : (c[0] > 0.2f ? RM_CODE_FLOOR_LIGHT : RM_CODE_FLOOR_DARK);
        }
        if (COUNT) c_sample = R;
        if (fp.max_iter > 0) live = (1u << R) - 1u;
    }

    float spill[RM_STACK_MAX - 1][R];
    float px_[R], py_[R], pz_[R], s[R];
    // ---------------- march: all live rays of the thread take one step per trip (wgsl:88-115) ----------------
    for (uint32_t it = 0; live != 0; ++it) {
        RM_FOR_R {  // pos = origin + direction * dist   wgsl:91
            px_[r] = A::mad(rdx[r], dist[r], ro[0]);
            py_[r] = A::mad(rdy[r], dist[r], ro[1]);
            pz_[r] = A::mad(rdz[r], dist[r], ro[2]);
        }
        RM_FOR_R asm volatile("" : "+f"(px_[r]), "+f"(py_[r]), "+f"(pz_[r]));
        map_evalR<EX, R, EO>(s_prog, fp.n_insn, fp.max_dist, px_, py_, pz_, s, spill);
        const bool last = it + 1 >= fp.max_iter;  // loop bound wgsl:90
        RM_FOR_R {
            const uint32_t bit = 1u << r;
            if (live & bit) {
                if (COUNT) ++c_step;
                if (s[r] < fp.min_dist) {  // wgsl:97: hit, dist stays at the hit distance
                    hit |= bit;
                    live &= ~bit;
                } else if (s[r] > fp.max_dist) {  // wgsl:109-111
                    live &= ~bit;
                } else {
                    dist[r] = A::add(dist[r], s[r]);  // wgsl:114
                    if (last) live &= ~bit;
                }
            }
        }
    }
    // ---------------- normals + diffuse term, one hit ray at a time (wgsl:98-105,138-143) ----------------
    const uint32_t hit0 = hit;
    while (hit != 0) {
        int rr[R / 4];
        float hx[R / 4], hy[R / 4], hz[R / 4];
#pragma unroll
        for (int j = 0; j < R / 4; ++j) {  // R/4 hit rays per trip: ray j's 4 taps fill slots 4j..4j+3
            rr[j] = hit ? __ffs((int)hit) - 1 : -1;
            hit &= hit - 1u;  // 0 stays 0
            float dx = 0.f, dy = 0.f, dz = 0.f, dd = 0.f;
            RM_FOR_R if (r == rr[j]) { dx = rdx[r]; dy = rdy[r]; dz = rdz[r]; dd = dist[r]; }
            hx[j] = A::mad(dx, dd, ro[0]); hy[j] = A::mad(dy, dd, ro[1]); hz[j] = A::mad(dz, dd, ro[2]);
            // tetrahedron taps wgsl:138-142: (+,-,-) (-,-,+) (-,+,-) (+,+,+), eps = 1e-4
            const float ep = RM_NORMAL_EPS, en = -RM_NORMAL_EPS;
            px_[4 * j + 0] = __fadd_rn(hx[j], ep); py_[4 * j + 0] = __fadd_rn(hy[j], en); pz_[4 * j + 0] = __fadd_rn(hz[j], en);
            px_[4 * j + 1] = __fadd_rn(hx[j], en); py_[4 * j + 1] = __fadd_rn(hy[j], en); pz_[4 * j + 1] = __fadd_rn(hz[j], ep);
            px_[4 * j + 2] = __fadd_rn(hx[j], en); py_[4 * j + 2] = __fadd_rn(hy[j], ep); pz_[4 * j + 2] = __fadd_rn(hz[j], en);
            px_[4 * j + 3] = __fadd_rn(hx[j], ep); py_[4 * j + 3] = __fadd_rn(hy[j], ep); pz_[4 * j + 3] = __fadd_rn(hz[j], ep);
        }
        RM_FOR_R asm volatile("" : "+f"(px_[r]), "+f"(py_[r]), "+f"(pz_[r]));
        map_evalR<EX, R, EO>(s_prog, fp.n_insn, fp.max_dist, px_, py_, pz_, s, spill);
#pragma unroll
        for (int j = 0; j < R / 4; ++j) {
            if (rr[j] >= 0) {
                const float c = wf_shade<EX>(s[4 * j], s[4 * j + 1], s[4 * j + 2], s[4 * j + 3], hx[j], hy[j], hz[j]);
                RM_FOR_R if (r == rr[j]) code[r] = c;
                if (COUNT) ++c_hit;
            }
        }
    }
    if (COUNT && inside) RM_FOR_R if (!(hit0 & (1u << r)) && code[r] != RM_CODE_SKY) ++c_floor;

    // ---------------- resolve: 16-sample mean in shader order (wgsl:44-45,68-75) ----------------
    // a pixel's threads hold consecutive sample ranges, so the running sum visits them in lane order
    float gm[R][3];
    RM_FOR_R code_to_gamma(code[r], gm[r]);
    float sr = 0.0f, sg = 0.0f, sb = 0.0f;
#pragma unroll
    for (int g = 0; g < TPP; ++g) {
        if ((int)grp == g) RM_FOR_R { sr = __fadd_rn(sr, gm[r][0]); sg = __fadd_rn(sg, gm[r][1]); sb = __fadd_rn(sb, gm[r][2]); }
        const int src = (lane / TPP) * TPP + g;
        sr = __shfl_sync(FULL, sr, src); sg = __shfl_sync(FULL, sg, src); sb = __shfl_sync(FULL, sb, src);
    }
    const uint32_t pix = unorm8(__fdiv_rn(sr, 16.0f)) | (unorm8(__fdiv_rn(sg, 16.0f)) << 8) |
                         (unorm8(__fdiv_rn(sb, 16.0f)) << 16) | 0xff000000u;
    // ---------------- store: each of the warp's ROWS rows is one 32-byte segment, two 128-bit stores ----------------
    const uint32_t x0 = tx * TILE_W;
    const int q4 = (lane & (2 * ROWS - 1)) * 4;  // first of the 4 pixels this lane gathers (used by lanes < 2*ROWS)
    uint4 v;
    v.x = __shfl_sync(FULL, pix, (q4 + 0) * TPP);
    v.y = __shfl_sync(FULL, pix, (q4 + 1) * TPP);
    v.z = __shfl_sync(FULL, pix, (q4 + 2) * TPP);
    v.w = __shfl_sync(FULL, pix, (q4 + 3) * TPP);
    const uint32_t row_q = (ty * 4u + (uint32_t)warp) * ROWS + (uint32_t)(q4 >> 3);
    if (x0 + TILE_W <= fp.W && (pitch & 15) == 0) {  // warp-uniform
        if (lane < 2 * ROWS && row_q < fp.nrows)
            *reinterpret_cast<uint4 *>(dst + (size_t)dst_row(fp, row_q) * pitch + (size_t)(x0 + (q4 & 7)) * 4) = v;
    } else if (inside && grp == 0) {
        *reinterpret_cast<uint32_t *>(dst + (size_t)dst_row(fp, ry) * pitch + (size_t)px * 4) = pix;
    }

    if (COUNT) {
        const uint32_t a = __reduce_add_sync(FULL, c_step), b = __reduce_add_sync(FULL, c_hit),
                       c = __reduce_add_sync(FULL, c_floor), d = __reduce_add_sync(FULL, c_sample);
        if (lane == 0) {
            atomicAdd(&counters[0], (unsigned long long)a);
            atomicAdd(&counters[1], (unsigned long long)b);
            atomicAdd(&counters[2], (unsigned long long)c);
            atomicAdd(&counters[3], (unsigned long long)d);
        }
    }
}

template <bool EX, int R, bool COUNT, bool EO>
cudaError_t launch_direct_t(const FrameParams &fp, const RmInsn *prog, uint8_t *dst, size_t pitch,
                            unsigned long long *counters, cudaStream_t st) {
    auto kern = k_direct<EX, R, COUNT, EO>;
    constexpr uint32_t tile_h = 4u * ((32u / (16u / R)) / 8u);
    const size_t smem = (size_t)fp.n_insn * sizeof(RmInsn);
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    const unsigned grid = ((fp.W + TILE_W - 1) / TILE_W) * ((fp.nrows + tile_h - 1) / tile_h);
    kern<<<grid, DIRECT_THREADS, smem, st>>>(fp, prog, dst, pitch, counters);
    return cudaGetLastError();
}

}  // namespace

cudaError_t launch_direct(bool exact, bool early_out, const FrameParams &fp, const RmInsn *prog_dev, uint8_t *dst, size_t pitch,
                          unsigned long long *counters, cudaStream_t st) {
    if (fp.nrows == 0 || fp.W == 0) return cudaSuccess;
    // Rays per thread: 4 for shallow programs (more threads, less lock-step waste), 8 for deep ones (fetch/dispatch
    // amortised over more rays).  Measured on B200 on truncated S2 chains at 1080p (tools/crossover.py, ms per frame,
    // R=4 / R=8): 1 instruction 0.72/0.85, 3: 1.38/1.47, 4: 1.80/1.82, 5: 2.10/2.09, 8: 3.21/3.14, 32: 12.39/11.87.
    // RM_DIRECT_R overrides.
    static const int kForce = [] {
        const char *s = std::getenv("RM_DIRECT_R");
        const int v = s ? std::atoi(s) : 0;
        return (v == 4 || v == 8) ? v : 0;
    }();
    const int kR = kForce ? kForce : (fp.n_insn <= 4 ? 4 : 8);
#define RM_DIRECT_DISPATCH2(EXV, RV, EOV)                                                      \
    (counters ? launch_direct_t<EXV, RV, true, EOV>(fp, prog_dev, dst, pitch, counters, st)     \
              : launch_direct_t<EXV, RV, false, EOV>(fp, prog_dev, dst, pitch, counters, st))
#define RM_DIRECT_DISPATCH(EXV, RV) (early_out ? RM_DIRECT_DISPATCH2(EXV, RV, true) : RM_DIRECT_DISPATCH2(EXV, RV, false))
    if (exact) return kR == 8 ? RM_DIRECT_DISPATCH(true, 8) : RM_DIRECT_DISPATCH(true, 4);
    return kR == 8 ? RM_DIRECT_DISPATCH(false, 8) : RM_DIRECT_DISPATCH(false, 4);
#undef RM_DIRECT_DISPATCH2
#undef RM_DIRECT_DISPATCH
}

}  // namespace rm
