// rm_direct_body.cuh — the body of the production kernel, shared by
//   * rm_direct.cu: compiled ahead of time around the decoded-program INTERPRETER (any program, no set-up cost), and
//   * rm_jit.cu:    compiled at run time (NVRTC) around map_scene emitted as STRAIGHT-LINE code for one program.
// It is everything of fs_main / ray_march (wgsl:33-115) except map_scene itself, which is the MAP argument: a functor
// `void operator()(const float (&px)[R], const float (&py)[R], const float (&pz)[R], float (&s)[R])` evaluating the scene
// at R points.  Must stay free of host headers (NVRTC has none).
//
// A thread owns R sample rays of one pixel (R = 2, 4 or 8) and marches them in lock-step (wgsl:87-115): the rays are a
// fraction of a pixel apart, so they leave the loop within a step or two of each other, and the warp's neighbouring
// pixels are coherent.  Rays that hit get their normal from the 4 tetrahedron taps (wgsl:138-142), which fill the same
// R slots of MAP, and the 16 samples of a pixel are summed in the shader's order by a shuffle chain over the pixel's
// 16/R threads (wgsl:68-75).  Per-ray arithmetic is operation for operation that of the other kernels.
#pragma once
#include "rm_device.cuh"

namespace rm {

#define RM_DIRECT_THREADS 128
#define RM_DIRECT_TILE_W 8
#define RM_FULL_MASK 0xffffffffu

// rows of the 8-wide pixel tile one 128-thread block covers
__host__ __device__ constexpr uint32_t direct_tile_rows(int R) { return 4u * (32u / (16u / (uint32_t)R)) / 8u; }
// The launch grid of the body: one block per 8 x direct_tile_rows(R) pixel tile, tile column in x and tile row in y (a
// linear grid cost every thread a 32-bit division to find its tile).  y holds up to 65535 tile rows = 131070 frame rows at
// 2 rays per thread.
__host__ __device__ inline dim3 direct_grid(const FrameParams &fp, uint32_t tile_h) {
    dim3 g;
    g.x = (fp.W + RM_DIRECT_TILE_W - 1) / RM_DIRECT_TILE_W;
    g.y = (fp.nrows + tile_h - 1) / tile_h;
    g.z = 1;
    return g;
}

// AOV (verification only): also writes every sample's class (0 sky, 1 floor, 2 object) and distance (hit distance, else
// the floor distance of wgsl:120) to aov_cls / aov_t[(row * W + x) * 16 + sample] — the same outputs, in the same
// layout, as the one-thread-per-sample verification kernel (rm_kernels.cu), but from THIS kernel's own march.
template <bool EX, int R, bool COUNT, bool AOV, class MAP>
__device__ __forceinline__ void direct_body(const FrameParams &fp, uint8_t *__restrict__ dst, size_t pitch,
                                            unsigned long long *__restrict__ counters, uint8_t *__restrict__ aov_cls,
                                            float *__restrict__ aov_t, MAP &map, uint32_t eval_flops = 0) {
    static_assert(R == 2 || R == 4 || R == 8, "2, 4 or 8 rays per thread");
    constexpr int TPP = 16 / R;           // threads per pixel
    constexpr int PPW = 32 / TPP;         // pixels per warp: 4, 8 or 16 consecutive pixels of the block's 8-wide tile
    constexpr uint32_t TILE_ROWS = direct_tile_rows(R);  // 2, 4 or 8
    using A = Arith<EX>;

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t tx = blockIdx.x, ty = blockIdx.y;  // direct_grid
    const uint32_t pb = (uint32_t)warp * PPW + (uint32_t)lane / TPP;  // pixel of the block, row-major over its 8-wide tile
    const uint32_t px = tx * RM_DIRECT_TILE_W + (pb & 7u), ry = ty * TILE_ROWS + (pb >> 3);
    const uint32_t grp = (uint32_t)lane % TPP;  // this thread holds samples grp*R .. grp*R+R-1 (sample = 4*i + j)
    const bool inside = px < fp.W && ry < fp.nrows;

    // The ray origin is the same for every ray of the frame (wgsl:39-40): read where it is used, it stays a uniform operand.
    const float ox = fp.ro4[0], oy = fp.ro4[1], oz = fp.ro4[2];
    float rdx[R], rdy[R], rdz[R], dist[R];
    uint32_t live = 0, hit = 0;  // bit r: ray r hit the scene; live: rays still marching (counting build only)
    uint32_t c_step = 0, c_hit = 0, c_floor = 0, c_sample = 0;
    RM_FOR_R { rdx[r] = rdy[r] = rdz[r] = 0.f; dist[r] = 0.f; }
    if (inside) {
        const uint32_t fy = frame_row(fp, ry);
        if constexpr (EX) {
            RM_FOR_R {
                float o[3], d[3];
                gen_ray<EX>(fp, px, fy, grp * R + (uint32_t)r, o, d);  // wgsl:39-62, one rounding per operation
                rdx[r] = d[0]; rdy[r] = d[1]; rdz[r] = d[2];
            }
        } else {
            const float fpx = (float)px + fp.ray_bx, fpy = (float)fy + fp.ray_by;
            RM_FOR_R {
                float d[3];
                gen_ray_affine(fp, fpx, fpy, grp * R + (uint32_t)r, d);  // the same rays from the per-frame affine form
                rdx[r] = d[0]; rdy[r] = d[1]; rdz[r] = d[2];
            }
        }
        if (COUNT) c_sample = R;
        live = (1u << R) - 1u;
    }

    float code[R];
    float t_miss[AOV ? R : 1];
    int cls_miss[AOV ? R : 1];
    uint32_t hit0 = 0;
    {
        float px_[R], py_[R], pz_[R], s[R];
        // ---------------- march: all live rays of the thread take one step per trip (wgsl:88-115) ----------------
        // (the loop bound of wgsl:90 is the same for every ray — they all start at trip 0 — so it bounds the loop itself:
        // rays still live when it runs out fall through to the floor / sky code exactly like aborted ones, Q5)
        // A stopped ray is not masked out of the later trips: its position is frozen, so MAP returns the same distance
        // for it every time, the two comparisons below come out the same way and its distance stays put (a fixed
        // point).  The loop therefore carries no live / hit masks: a ray has hit iff the distance of the LAST trip is
        // under min_dist (a ray still marching when the bound runs out was not), and the thread goes on while any of
        // its rays advanced.  Per ray and trip: two compares and one predicated add, no divergent branch.
        bool go = inside;
        RM_FOR_R s[r] = 3.0e38f;
        for (uint32_t it = 0; go && it < fp.max_iter; ++it) {
            RM_FOR_R {  // pos = origin + direction * dist   wgsl:91
                px_[r] = A::mad(rdx[r], dist[r], ox);
                py_[r] = A::mad(rdy[r], dist[r], oy);
                pz_[r] = A::mad(rdz[r], dist[r], oz);
            }
            RM_FOR_R RM_OPAQUE_F3(px_[r], py_[r], pz_[r]);
            map(px_, py_, pz_, s);
            go = false;
            RM_FOR_R {
                // wgsl:97 hit, dist stays at the hit distance; wgsl:109-111 gone past max_dist; else wgsl:114 advance
                const bool stop = (s[r] < fp.min_dist) | (s[r] > fp.max_dist);
                if (COUNT) {
                    c_step += (live >> r) & 1u;
                    live &= stop ? ~(1u << r) : ~0u;
                }
                dist[r] = stop ? dist[r] : A::add(dist[r], s[r]);  // wgsl:114
                go |= !stop;
            }
        }
        RM_FOR_R hit |= (s[r] < fp.min_dist) ? (1u << r) : 0u;
        // ---------------- normals + diffuse term of the rays that hit (wgsl:98-105,138-143) ----------------
        hit0 = hit;
        RM_FOR_R code[r] = RM_CODE_SKY;
        // floor or sky for the rays that never hit (wgsl:117-130): after the march so that nothing of it is live across the
        // loop; always with exact roundings (one division per missed ray).  With 2 rays per thread it runs BEFORE the taps
        // (the ray directions are then dead while MAP runs again), else after them.
        auto shade_missed = [&]() {
            uint32_t tx_m, ty_m;
            RM_REREAD_CTAID(tx_m, ty_m);  // `inside` again, from scratch: nothing of it lives across the march
            const bool inside_m = tx_m * RM_DIRECT_TILE_W + (pb & 7u) < fp.W && ty_m * TILE_ROWS + (pb >> 3) < fp.nrows;
            RM_FOR_R {
                if constexpr (AOV) { t_miss[r] = 0.f; cls_miss[r] = 0; }
                if (inside_m && !((hit0 >> r) & 1u)) {
                    const float o[3] = {ox, oy, oz}, d[3] = {rdx[r], rdy[r], rdz[r]};
                    float t;
                    int cls;
                    code[r] = shade_miss_code(o, d, t, cls);
                    if (COUNT && cls != 0) ++c_floor;
                    if constexpr (AOV) { t_miss[r] = t; cls_miss[r] = cls; }
                }
            }
        };
        if constexpr (R == 2) shade_missed();
        const float ep = RM_NORMAL_EPS, en = -RM_NORMAL_EPS;
        if constexpr (R == 2) {
            // Both rays' hit positions first: the directions and distances are then dead, and the tap loop below carries
            // only what it needs across its two inlined evaluations of MAP.  A hit ray's 4 taps take two passes of the
            // 2-wide map: taps 0,1 then taps 2,3.
            const float h0x = A::mad(rdx[0], dist[0], ox), h0y = A::mad(rdy[0], dist[0], oy), h0z = A::mad(rdz[0], dist[0], oz);
            const float h1x = A::mad(rdx[1], dist[1], ox), h1y = A::mad(rdy[1], dist[1], oy), h1z = A::mad(rdz[1], dist[1], oz);
            while (hit != 0) {
                const int rr = __ffs((int)hit) - 1;
                hit &= hit - 1u;
                const float hx = rr ? h1x : h0x, hy = rr ? h1y : h0y, hz = rr ? h1z : h0z;
                float a0 = 0.f, a1 = 0.f;
                float c = 0.f;
    #pragma unroll 1
                for (int pass = 0; pass < 2; ++pass) {  // taps wgsl:138-142: (+,-,-) (-,-,+) | (-,+,-) (+,+,+)
                    px_[0] = __fadd_rn(hx, pass ? en : ep); py_[0] = __fadd_rn(hy, pass ? ep : en); pz_[0] = __fadd_rn(hz, en);
                    px_[1] = __fadd_rn(hx, pass ? ep : en); py_[1] = __fadd_rn(hy, pass ? ep : en); pz_[1] = __fadd_rn(hz, ep);
                    map(px_, py_, pz_, s);
                    if (pass == 0) { a0 = s[0]; a1 = s[1]; }
                    else c = wf_shade<EX>(a0, a1, s[0], s[1], hx, hy, hz);
                }
                if (rr) code[1] = c; else code[0] = c;
                if (COUNT) ++c_hit;
            }
        } else {
            while (hit != 0) {
                int rr[(R + 3) / 4];
                float hx[(R + 3) / 4], hy[(R + 3) / 4], hz[(R + 3) / 4];
    #pragma unroll
                for (int j = 0; j < R / 4; ++j) {  // R/4 hit rays per trip: ray j's 4 taps fill slots 4j..4j+3
                    rr[j] = hit ? __ffs((int)hit) - 1 : -1;
                    hit &= hit - 1u;  // 0 stays 0
                    float dx = 0.f, dy = 0.f, dz = 0.f, dd = 0.f;
                    RM_FOR_R if (r == rr[j]) { dx = rdx[r]; dy = rdy[r]; dz = rdz[r]; dd = dist[r]; }
                    hx[j] = A::mad(dx, dd, ox); hy[j] = A::mad(dy, dd, oy); hz[j] = A::mad(dz, dd, oz);
                    // tetrahedron taps wgsl:138-142: (+,-,-) (-,-,+) (-,+,-) (+,+,+), eps = 1e-4
                    px_[4 * j + 0] = __fadd_rn(hx[j], ep); py_[4 * j + 0] = __fadd_rn(hy[j], en); pz_[4 * j + 0] = __fadd_rn(hz[j], en);
                    px_[4 * j + 1] = __fadd_rn(hx[j], en); py_[4 * j + 1] = __fadd_rn(hy[j], en); pz_[4 * j + 1] = __fadd_rn(hz[j], ep);
                    px_[4 * j + 2] = __fadd_rn(hx[j], en); py_[4 * j + 2] = __fadd_rn(hy[j], ep); pz_[4 * j + 2] = __fadd_rn(hz[j], en);
                    px_[4 * j + 3] = __fadd_rn(hx[j], ep); py_[4 * j + 3] = __fadd_rn(hy[j], ep); pz_[4 * j + 3] = __fadd_rn(hz[j], ep);
                }
                RM_FOR_R RM_OPAQUE_F3(px_[r], py_[r], pz_[r]);
                map(px_, py_, pz_, s);
    #pragma unroll
                for (int j = 0; j < R / 4; ++j) {
                    if (rr[j] >= 0) {
                        const float c = wf_shade<EX>(s[4 * j], s[4 * j + 1], s[4 * j + 2], s[4 * j + 3], hx[j], hy[j], hz[j]);
                        RM_FOR_R if (r == rr[j]) code[r] = c;
                        if (COUNT) ++c_hit;
                    }
                }
            }
        }
        if constexpr (R != 2) shade_missed();
    }
    // the tile's coordinates once more for the outputs (see RM_REREAD_CTAID)
    uint32_t tx_o, ty_o;
    RM_REREAD_CTAID(tx_o, ty_o);
    const uint32_t px_o = tx_o * RM_DIRECT_TILE_W + (pb & 7u), ry_o = ty_o * TILE_ROWS + (pb >> 3);
    const bool inside_o = px_o < fp.W && ry_o < fp.nrows;
    if constexpr (AOV) if (inside_o) {
        const size_t base = ((size_t)ry_o * fp.W + px_o) * 16 + grp * R;
        RM_FOR_R {
            const bool h = (hit0 >> r) & 1u;
            aov_cls[base + r] = (uint8_t)(h ? 2 : cls_miss[r]);
            aov_t[base + r] = h ? dist[r] : t_miss[r];
        }
    }

    // ---------------- resolve: 16-sample mean in shader order (wgsl:44-45,68-75) ----------------
    // a pixel's threads hold consecutive sample ranges, so the running sum visits them in lane order
    float gm[R][3];
    RM_FOR_R code_to_gamma_t<EX>(code[r], gm[r]);
    float sr = 0.0f, sg = 0.0f, sb = 0.0f;
    if constexpr (EX) {
#pragma unroll
        for (int g = 0; g < TPP; ++g) {
            if ((int)grp == g) RM_FOR_R { sr = __fadd_rn(sr, gm[r][0]); sg = __fadd_rn(sg, gm[r][1]); sb = __fadd_rn(sb, gm[r][2]); }
            const int src = (lane / TPP) * TPP + g;
            sr = __shfl_sync(RM_FULL_MASK, sr, src); sg = __shfl_sync(RM_FULL_MASK, sg, src); sb = __shfl_sync(RM_FULL_MASK, sb, src);
        }
    } else {
        // fast mode: the same 16 terms as a tree — each thread its own R, then a butterfly over the pixel's lanes (every
        // lane ends with the same bits: a + b and b + a).  Differs from the sequential sum by roundings of ~1e-7 relative.
        RM_FOR_R { sr += gm[r][0]; sg += gm[r][1]; sb += gm[r][2]; }
#pragma unroll
        for (int d = 1; d < TPP; d <<= 1) {
            sr += __shfl_xor_sync(RM_FULL_MASK, sr, d); sg += __shfl_xor_sync(RM_FULL_MASK, sg, d); sb += __shfl_xor_sync(RM_FULL_MASK, sb, d);
        }
    }
    const uint32_t pix = pack_mean16(sr, sg, sb);
    // ---------------- store: every 4 consecutive pixels of a row go out as one 128-bit store ----------------
    const uint32_t x0 = tx_o * RM_DIRECT_TILE_W;
    constexpr int QUADS = PPW / 4;               // 4-pixel groups per warp
    const int q4 = (lane & (QUADS - 1)) * 4;     // first of the 4 pixels this lane gathers (used by lanes < QUADS)
    uint4 v;
    v.x = __shfl_sync(RM_FULL_MASK, pix, (q4 + 0) * TPP);
    v.y = __shfl_sync(RM_FULL_MASK, pix, (q4 + 1) * TPP);
    v.z = __shfl_sync(RM_FULL_MASK, pix, (q4 + 2) * TPP);
    v.w = __shfl_sync(RM_FULL_MASK, pix, (q4 + 3) * TPP);
    const uint32_t pb_q = (uint32_t)warp * PPW + (uint32_t)q4;
    const uint32_t row_q = ty_o * TILE_ROWS + (pb_q >> 3);
    if (x0 + RM_DIRECT_TILE_W <= fp.W && (pitch & 15) == 0) {  // warp-uniform
        if (lane < QUADS && row_q < fp.nrows)
            *reinterpret_cast<uint4 *>(dst + (size_t)dst_row(fp, row_q) * pitch + (size_t)(x0 + (pb_q & 7u)) * 4) = v;
    } else if (inside_o && grp == 0) {
        *reinterpret_cast<uint32_t *>(dst + (size_t)dst_row(fp, ry_o) * pitch + (size_t)px_o * 4) = pix;
    }

    if (COUNT) {
        const uint32_t a = __reduce_add_sync(RM_FULL_MASK, c_step), b = __reduce_add_sync(RM_FULL_MASK, c_hit),
                       c = __reduce_add_sync(RM_FULL_MASK, c_floor), d = __reduce_add_sync(RM_FULL_MASK, c_sample);
        if (lane == 0) {
            atomicAdd(&counters[0], (unsigned long long)a);
            atomicAdd(&counters[1], (unsigned long long)b);
            atomicAdd(&counters[2], (unsigned long long)c);
            atomicAdd(&counters[3], (unsigned long long)d);
            // culled paints (rm_cull.cu): algorithmic flops of the map_scene evaluations as EXECUTED with this block's program
            if (eval_flops) atomicAdd(&counters[4], ((unsigned long long)a + 4ull * b) * eval_flops);
        }
    }
}

}  // namespace rm
