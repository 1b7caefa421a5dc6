// rm_cull.cu — opt-in, never the headline (rm_set_cull, bench.py --cull on): per-tile conservative culling of the decoded
// program, exact by construction.
//
// For one fused instruction  acc = op(acc, prim(p))  the accumulator is unchanged BIT FOR BIT at a point p when
//   union / smooth union:  prim(p) - acc(p) >= k          (k = 0 for the plain union: e = max(k - |acc - v|, 0) == 0, min = acc)
//   subtraction:           prim(p) + acc(p) >= 0          (max(acc, -v) == acc)
//   intersection:          acc(p) - prim(p) >= 0          (max(acc, v) == acc)
// (the same facts the per-evaluation early-out of rm_wavefront_common.cuh uses).  Sphere and box distances are 1-Lipschitz
// and so is every accumulator value (min, max and the polynomial smooth-min of 1-Lipschitz functions are), so if the
// margin above, evaluated at ONE point c, exceeds 2*delta (+ slack), the instruction is a no-op at every point within delta
// of c.  One thread per pixel tile walks the tile's centre ray from t = 0 to beyond the last position any ray of the frame
// can visit, in intervals [t, t + w]: the ball of radius  w + (t + w)*chi + 2e-4  around the centre-ray point at t covers
// every point a ray of the tile can be at for a parameter in that interval (chi = the largest chord between the centre
// direction and a direction of the tile, from its four corners; 2e-4 = the normal taps' offsets).  An instruction whose
// margin clears the bound at EVERY sample is dropped from that tile's program; everything else (first / push / pop
// instructions, and any instruction that fails once) stays.  By induction over the instruction stream the accumulator at
// every visited point is then the same bits as with the full program: image, AOVs and work counters are identical
// (tests/test_gpu_cull.py), only the number of instructions executed per evaluation differs — which is why this mode is
// reported with EXECUTED flops beside the algorithmic ones and is never the roofline headline.
#include "rm_cull.h"

#include <mutex>

#include "rm_program.h"

namespace rm {

namespace {

constexpr int kMaxSamples = 8192;
constexpr float kMaxWidth = 0.25f;
constexpr float kSlack = 1e-3f;  // float evaluation error of the margins (values of a few units: ~1e-6) with a wide berth

__device__ __forceinline__ void affine_dir(const FrameParams &fp, float X, float Y, float d[3]) {
    const float dx = fmaf(Y, fp.ray_cy[0], fmaf(X, fp.ray_cx[0], fp.ray_c0[0]));
    const float dy = fmaf(Y, fp.ray_cy[1], fmaf(X, fp.ray_cx[1], fp.ray_c0[1]));
    const float dz = fmaf(Y, fp.ray_cy[2], fmaf(X, fp.ray_cx[2], fp.ray_c0[2]));
    const float dw = fmaf(Y, fp.ray_cy[3], fmaf(X, fp.ray_cx[3], fp.ray_c0[3]));
    const float inv = rsqrtf(fmaf(dw, dw, fmaf(dz, dz, fmaf(dy, dy, dx * dx))));
    d[0] = dx * inv; d[1] = dy * inv; d[2] = dz * inv;
}

// One thread per tile.  prog: the decoded program the paint will run (exact or fast form), n_insn <= 64.
__global__ void __launch_bounds__(64) k_cull_masks(const __grid_constant__ FrameParams fp, const RmInsn *__restrict__ prog, uint32_t tile_h,
                                                   uint32_t tiles_x, uint32_t n_tiles, float t_end, unsigned long long *__restrict__ masks) {
    extern __shared__ RmInsn s_prog[];
    for (uint32_t i = threadIdx.x; i < fp.n_insn; i += blockDim.x) s_prog[i] = prog[i];
    __syncthreads();
    const uint32_t tile = blockIdx.x * blockDim.x + threadIdx.x;
    if (tile >= n_tiles) return;
    const uint32_t n = fp.n_insn;
    const unsigned long long all = n >= 64 ? ~0ull : ((1ull << n) - 1ull);
    const uint32_t tx = tile % tiles_x, ty = tile / tiles_x;
    // pixel rectangle of the tile in FRAME coordinates (the tile's rows are contiguous frame rows: launch_cull_masks checks)
    const float x0 = (float)(tx * 8u), x1 = fminf((float)(tx * 8u + 8u), (float)fp.W);
    const uint32_t ry0 = ty * tile_h, ry1 = min(ty * tile_h + tile_h, fp.nrows);
    const float y0 = (float)frame_row(fp, ry0), y1 = (float)(frame_row(fp, ry1 - 1u) + 1u);
    // sample positions span [px + 1/8, px + 7/8] per pixel (wgsl:52-53): corners of the pixel rectangle bound them; the
    // affine form wants X = px + ray_bx + i*ray_gx: pixel-edge coordinates map to X = edge + (ray_bx - 0.5 + 1.5*ray_gx) = edge
    // (a viewport_extent smaller than the frame spreads the samples beyond their pixel: pad by the excess)
    const float padx = fmaxf(1.5f * fabsf(fp.ray_gx) - 0.5f, 0.f), pady = fmaxf(1.5f * fabsf(fp.ray_gy) - 0.5f, 0.f);
    float dc[3] = {0.f, 0.f, 0.f}, dk[4][3];
    const float xs[2] = {x0 - padx, x1 + padx}, ys[2] = {y0 - pady, y1 + pady};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        affine_dir(fp, xs[k & 1], ys[k >> 1], dk[k]);
        dc[0] += dk[k][0]; dc[1] += dk[k][1]; dc[2] += dk[k][2];
    }
    {
        const float inv = rsqrtf(dc[0] * dc[0] + dc[1] * dc[1] + dc[2] * dc[2]);
        dc[0] *= inv; dc[1] *= inv; dc[2] *= inv;
    }
    float chi = 0.f;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float ex = dk[k][0] - dc[0], ey = dk[k][1] - dc[1], ez = dk[k][2] - dc[2];
        chi = fmaxf(chi, sqrtf(ex * ex + ey * ey + ez * ez));
    }
    chi = chi * 1.05f + 4e-6f;  // exact-mode rays differ from the affine form by a few 1e-7
    if (!(chi < 0.5f)) { masks[tile] = all; return; }  // degenerate camera: no culling

    const float ox = fp.ro4[0], oy = fp.ro4[1], oz = fp.ro4[2];
    unsigned long long active = 0ull;
    float margin[64];
    float t = 0.f;
    int samples = 0;
    while (t <= t_end) {
        if (++samples > kMaxSamples) { active = all; break; }
        const float px = fmaf(dc[0], t, ox), py = fmaf(dc[1], t, oy), pz = fmaf(dc[2], t, oz);
        // the program at this point, sequentially, with the FULL instruction set (margins refer to the true accumulator)
        float acc = 0.f;
        float spill[RM_STACK_MAX - 1];
        for (uint32_t i = 0; i < n; ++i) {
            const RmInsn in = s_prog[i];
            const uint32_t fl = in.flags;
            margin[i] = -1.f;  // never dropped unless a fused leaf instruction proves otherwise
            if (fl & FLAG_POP) {
                const float a = spill[in.slot], b = acc;
                if (fl & FLAG_SMOOTH) { const float e = fmaxf(in.k - fabsf(a - b), 0.f); acc = fminf(a, b) - e * e * (0.25f / in.k); }
                else if (fl & FLAG_UNION) acc = fminf(a, b);
                else if (fl & FLAG_SUB) acc = fmaxf(a, -b);
                else acc = fmaxf(a, b);
                continue;
            }
            float v;
            {
                const float dx = px - in.cx, dy = py - in.cy, dz = pz - in.cz;
                if (fl & FLAG_BOX) {
                    const float qx = fabsf(dx) - in.r0, qy = fabsf(dy) - in.r1, qz = fabsf(dz) - in.r2;
                    const float mx = fmaxf(qx, 0.f), my = fmaxf(qy, 0.f), mz = fmaxf(qz, 0.f);
                    v = sqrtf(mx * mx + my * my + mz * mz) + fminf(fmaxf(qx, fmaxf(qy, qz)), 0.f);
                } else {
                    v = sqrtf(dx * dx + dy * dy + dz * dz) - in.r0;
                }
            }
            if (fl & FLAG_SMOOTH) {
                margin[i] = (v - acc) - in.k;
                const float e = fmaxf(in.k - fabsf(acc - v), 0.f);
                acc = fminf(acc, v) - e * e * (0.25f / in.k);
            } else if (fl & FLAG_UNION) {
                margin[i] = v - acc;
                acc = fminf(acc, v);
            } else if (fl & FLAG_SUB) {
                margin[i] = v + acc;
                acc = fmaxf(acc, -v);
            } else if (fl & FLAG_INTER) {
                margin[i] = acc - v;
                acc = fmaxf(acc, v);
            } else {
                if (fl & FLAG_PUSH) spill[in.slot] = acc;
                acc = v;
            }
        }
        // interval width: fine near surfaces, a quarter of the scene distance further out, but never wide — a margin has
        // to clear TWICE the ball radius, and far from the scene the leaves' distances differ by no more than the scene is
        // deep, so wide intervals there would keep every instruction alive for the (many) far steps of the escaping rays
        const float w = fminf(fmaxf(0.25f * fabsf(acc), fmaxf(0.02f, t * chi)), kMaxWidth);
        const float bound = 2.f * (w + (t + w) * chi + 2e-4f) + kSlack;
        for (uint32_t i = 0; i < n; ++i)
            if (!(margin[i] >= bound)) active |= 1ull << i;
        if (active == all) break;
        t += w;
    }
    masks[tile] = active;
}

}  // namespace

cudaError_t launch_cull_masks(const FrameParams &fp, const RmInsn *prog_dev, uint32_t tile_h, float t_end, unsigned long long *masks_dev,
                              cudaStream_t st) {
    if (fp.nrows == 0 || fp.W == 0) return cudaSuccess;
    const uint32_t tiles_x = (fp.W + 7u) / 8u, tiles_y = (fp.nrows + tile_h - 1u) / tile_h, n_tiles = tiles_x * tiles_y;
    const unsigned blocks = (n_tiles + 63u) / 64u;
    k_cull_masks<<<blocks, 64, (size_t)fp.n_insn * sizeof(RmInsn), st>>>(fp, prog_dev, tile_h, tiles_x, n_tiles, t_end, masks_dev);
    return cudaGetLastError();
}

// ---- allocation-free variant (rm_paint_culled_static; no counterpart in the reference)
// -----------------------------------------------------------------------------------------------------------
// The masks of rm_paint_culled_static live in this file-scope device array instead of a cudaMalloc'ed buffer: callers that
// cannot allocate (CUDA graph capture, tight serving loops) get the culled paint without any allocation.  Every GPU that
// loads the library holds the RM_CULL_STATIC_MAX_TILES * 8 bytes (2 MiB) until exit.  In HOST code the array's name is only
// a stand-in for the symbol, not a device address: the address comes from cudaGetSymbolAddress, once per device with that
// device current, and that pointer is what the kernels are given.  The pre-pass writes the mask of every tile the paint
// reads (nothing has to start at zero); the array keeps the last call's contents, and calls that overlap on one GPU race.
static __device__ __align__(256) unsigned long long g_cull_static_masks[RM_CULL_STATIC_MAX_TILES];

cudaError_t cull_static_masks(int device, unsigned long long **ptr) {
    static std::mutex mu;
    static unsigned long long *ptrs[64] = {};
    if (!ptr || device < 0 || device >= 64) return cudaErrorInvalidValue;
    std::lock_guard<std::mutex> lock(mu);
    if (!ptrs[device]) {
        int cur = -1;
        cudaError_t e = cudaGetDevice(&cur);
        if (e != cudaSuccess) return e;
        if (cur != device) return cudaErrorInvalidDevice;  // the symbol's address belongs to the device that is current
        void *p = nullptr;
        e = cudaGetSymbolAddress(&p, g_cull_static_masks);
        if (e != cudaSuccess) return e;
        ptrs[device] = static_cast<unsigned long long *>(p);
    }
    *ptr = ptrs[device];
    return cudaSuccess;
}

// The pre-pass of launch_cull_masks into the static array of `device` (which must be current): same kernel, the masks go
// to the symbol's device address; *masks_out receives that address for the paint kernel.
cudaError_t launch_cull_masks_static(int device, const FrameParams &fp, const RmInsn *prog_dev, uint32_t tile_h, float t_end,
                                     const unsigned long long **masks_out, cudaStream_t st) {
    const size_t n_tiles = (size_t)((fp.W + 7u) / 8u) * ((fp.nrows + tile_h - 1u) / tile_h);
    if (n_tiles > RM_CULL_STATIC_MAX_TILES) return cudaErrorInvalidValue;
    unsigned long long *masks = nullptr;
    cudaError_t e = cull_static_masks(device, &masks);
    if (e != cudaSuccess) return e;
    *masks_out = masks;
    return launch_cull_masks(fp, prog_dev, tile_h, t_end, masks, st);
}

}  // namespace rm
