// rm_wavefront2.cu — second-generation persistent wavefront kernel of the render pass.
//
// Why it exists (profiles/r1_s1_NOTES.md): in rm_wavefront.cu the interpreter body already runs at the
// zero-overhead arithmetic floor, but 34 % of the time goes to per-iteration bookkeeping that issues
// badly — a 5-state machine per ray slot with divergent branches, one shared-memory atomic per finished
// ray (8 WARPSYNCs per iteration), state selects in the evaluation-point setup, divergent shading.
// Here a ray slot only ever MARCHES (wgsl:87-115):
//   * the per-slot update is branch-free (predicated selects); finished rays are counted in packed
//     per-lane counters and folded into the tile's `remaining` count with two warp reductions;
//   * a ray that hits (wgsl:97) leaves its slot at once: its hit position is pushed to a per-warp
//     hit queue.  When the queue holds a full batch (32*R/4 rays) the warp spends one iteration on
//     normals: every lane takes R/4 queued rays and evaluates their 4 tetrahedron taps (wgsl:138-142)
//     in its R slots through the SAME interpreter body, then shades them in-lane (wgsl:98-105).
// Arithmetic per ray is unchanged operation for operation, so images, classes and counters are
// bit-identical to rm_wavefront.cu and to the one-thread-per-sample kernel in both modes.
#include "rm_wavefront.h"
#include "rm_wavefront_common.cuh"

#include <cstdlib>

namespace rm {

namespace {

constexpr int HQ_CAP = 192;  // hit-queue entries per warp; a hit that finds it full retries next iteration

template <int R>
struct WarpShared2 {
    float codes[WF_K][TILE_RAYS];  // per-sample result codes; pre-set to the ray's miss code
    float4 ring[RING_N];           // generated rays waiting for a slot: dir.xyz + meta bits
    float4 hitq[HQ_CAP];           // rays that hit, waiting for their normal: hit position + (q | buffer) bits
    int remaining[WF_K];
    int tile[WF_K];  // tile id, or -1 when the buffer is free
};

constexpr uint32_t M_QB = M_Q | (3u << M_BUF_SHIFT);  // sample slot + tile buffer of a ray

template <bool EX, int R, bool COUNT>
__global__ void __launch_bounds__(WF_WARPS * 32, (R >= 8 ? 4 : 5))
k_wavefront2(const __grid_constant__ FrameParams fp, const RmInsn *__restrict__ prog_g, uint8_t *__restrict__ dst,
             size_t pitch, unsigned int *__restrict__ tile_counter, unsigned long long *__restrict__ counters) {
    static_assert(R % 4 == 0, "a lane evaluates the 4 normal taps of R/4 rays in its R slots");
    constexpr int RPL = R / 4;     // queued rays per lane in a normal pass
    constexpr int NB = 32 * RPL;   // rays per normal pass
    using A = Arith<EX>;
    extern __shared__ float4 smem_f4[];
    float4 *s_prog = smem_f4;
    const uint32_t nq = fp.n_insn * 3;
    WarpShared2<R> *s_warps = reinterpret_cast<WarpShared2<R> *>(smem_f4 + nq);
    for (uint32_t i = threadIdx.x; i < nq; i += blockDim.x) s_prog[i] = reinterpret_cast<const float4 *>(prog_g)[i];
    const int lane = threadIdx.x & 31;
    WarpShared2<R> &ws = s_warps[threadIdx.x >> 5];
    if (lane < WF_K) {
        ws.remaining[lane] = 0;
        ws.tile[lane] = -1;
    }
    __syncthreads();

    const uint32_t tiles_x = (fp.W + TILE_W - 1) / TILE_W;
    const uint32_t tiles_y = (fp.nrows + TILE_H - 1) / TILE_H;
    const uint32_t n_tiles = tiles_x * tiles_y;
    const unsigned lt_mask = (1u << lane) - 1u;

    // ray origin = inv_view * (0,0,0,1)  (wgsl:39-40) — identical for every ray of the frame
    float ro[3];
    {
        ro[0] = fp.ro4[0]; ro[1] = fp.ro4[1]; ro[2] = fp.ro4[2];
    }

    float rdx[R], rdy[R], rdz[R], dist[R];
    uint32_t meta[R];
    float spill[RM_STACK_MAX - 1][R];
    RM_FOR_R { meta[r] = 0; dist[r] = 0.f; rdx[r] = rdy[r] = rdz[r] = 0.f; }

    int cur_buf = 0, pool_next = TILE_RAYS, ring_head = 0, ring_count = 0, hq_count = 0;
    uint32_t cur_tile = 0;
    bool exhausted = false;
    uint32_t c_step = 0, c_hit = 0, c_floor = 0, c_sample = 0;

    for (;;) {
        bool starved = false;  // wanted to open a tile, but every buffer is still waiting for results
        bool check = false;    // some tile's `remaining` changed this iteration
        // ---------------- ray generation: 32 rays per pass, one per lane, into the ready ring -------
        // (full-lane work; the miss colour depends only on the ray, so it is computed here too and
        // parked as the sample's default result — a ray that misses later has nothing left to do)
        if (ring_count <= RING_N - 32) {  // warp-uniform
            if (pool_next >= TILE_RAYS && !exhausted) {  // open the next tile if a buffer is free
                int b = -1;
#pragma unroll
                for (int k = WF_K - 1; k >= 0; --k) if (ws.tile[k] < 0) b = k;
                if (b >= 0) {
                    unsigned t = 0;
                    if (lane == 0) t = atomicAdd(tile_counter, 1u);
                    t = __shfl_sync(FULL, t, 0);
                    if (t >= n_tiles) {
                        exhausted = true;
                    } else {
                        cur_buf = b; cur_tile = t; pool_next = 0;
                        if (lane == 0) { ws.tile[b] = (int)t; ws.remaining[b] = TILE_RAYS; }
                        __syncwarp();
                    }
                } else {
                    starved = true;
                }
            }
            if (pool_next < TILE_RAYS) {
                const uint32_t q = (uint32_t)pool_next + lane;  // sample = pool_next/32, pixel = lane
                const uint32_t tx = cur_tile % tiles_x, ty = cur_tile / tiles_x;
                const uint32_t px = tx * TILE_W + (lane & 7u), ry = ty * TILE_H + (lane >> 3);
                const bool inside = px < fp.W && ry < fp.nrows;
                float code = RM_CODE_SKY;
                float o[3], d[3] = {0.f, 0.f, 0.f};
                if (inside) {
                    gen_ray<EX>(fp, px, frame_row(fp, ry), q >> 5, o, d);
                    float c[3], t;
                    const int cls = shade_miss(ro, d, c, t);  // wgsl:117-130
                    code = cls == 0 ? RM_CODE_SKY : (c[0] > 0.2f ? RM_CODE_FLOOR_LIGHT : RM_CODE_FLOOR_DARK);
                    if (COUNT) ++c_sample;
                }
                ws.codes[cur_buf][q] = code;
                const bool live = inside && fp.max_iter > 0;
                const unsigned lm = __ballot_sync(FULL, live);
                if (live) {
                    const int pos = (ring_head + ring_count + __popc(lm & lt_mask)) & (RING_N - 1);
                    ws.ring[pos] = make_float4(d[0], d[1], d[2],
                                               __uint_as_float(q | ((uint32_t)cur_buf << M_BUF_SHIFT) | M_LIVE));
                } else if (COUNT && code != RM_CODE_SKY) {
                    ++c_floor;
                }
                if (lm != FULL) {  // rays with nothing to march (outside the band, or max_iter == 0)
                    if (lane == 0) ws.remaining[cur_buf] -= 32 - __popc(lm);
                    check = true;
                }
                ring_count += __popc(lm);
                pool_next += 32;
                __syncwarp();
            }
        }
        // ---------------- refill empty slots from the ready ring (ballot compaction) ----------------
        RM_FOR_R {
            const bool empty = !(meta[r] & M_LIVE);
            const unsigned em = __ballot_sync(FULL, empty);
            if (em != 0 && ring_count > 0) {  // warp-uniform
                const int rank = __popc(em & lt_mask);
                const int take = min(__popc(em), ring_count);
                if (empty && rank < take) {
                    const float4 e = ws.ring[(ring_head + rank) & (RING_N - 1)];
                    rdx[r] = e.x; rdy[r] = e.y; rdz[r] = e.z;
                    dist[r] = 0.0f;
                    meta[r] = __float_as_uint(e.w);
                }
                ring_head = (ring_head + take) & (RING_N - 1);
                ring_count -= take;
            }
        }
        bool live_mine = false;
        RM_FOR_R live_mine |= (meta[r] & M_LIVE) != 0;
        const bool live_any = __any_sync(FULL, live_mine);
        // a normal pass when a full batch is queued, or when the slots have nothing better to do
        const bool normal = hq_count >= NB || (hq_count > 0 && (!live_any || starved));
        uint32_t dec_lo = 0, dec_hi = 0;  // finished rays per tile buffer, 16-bit fields: lo = buffers 0,1; hi = 2,3
        if (live_any || hq_count > 0) {  // warp-uniform: there is something to evaluate
            // ---------------- evaluation points ----------------
            float px[R], py[R], pz[R], s[R];
            const int n_take = min(hq_count, NB);
            const int hq_base = hq_count - n_take;  // normal passes pop from the top of the queue
            if (!normal) {
                RM_FOR_R {  // pos = origin + direction * dist   wgsl:91
                    px[r] = A::mad(rdx[r], dist[r], ro[0]);
                    py[r] = A::mad(rdy[r], dist[r], ro[1]);
                    pz[r] = A::mad(rdz[r], dist[r], ro[2]);
                }
            } else {
#pragma unroll
                for (int j = 0; j < RPL; ++j) {
                    const int e = lane * RPL + j;
                    float4 h = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (e < n_take) h = ws.hitq[hq_base + e];
                    // tetrahedron taps wgsl:138-142: (+,-,-) (-,-,+) (-,+,-) (+,+,+), eps = 1e-4
                    const float ep = RM_NORMAL_EPS, en = -RM_NORMAL_EPS;
                    px[4 * j + 0] = __fadd_rn(h.x, ep); py[4 * j + 0] = __fadd_rn(h.y, en); pz[4 * j + 0] = __fadd_rn(h.z, en);
                    px[4 * j + 1] = __fadd_rn(h.x, en); py[4 * j + 1] = __fadd_rn(h.y, en); pz[4 * j + 1] = __fadd_rn(h.z, ep);
                    px[4 * j + 2] = __fadd_rn(h.x, en); py[4 * j + 2] = __fadd_rn(h.y, ep); pz[4 * j + 2] = __fadd_rn(h.z, en);
                    px[4 * j + 3] = __fadd_rn(h.x, ep); py[4 * j + 3] = __fadd_rn(h.y, ep); pz[4 * j + 3] = __fadd_rn(h.z, ep);
                }
            }
            // keep the points in registers: without this ptxas rematerialises the setup inside every
            // interpreted instruction
            RM_FOR_R asm volatile("" : "+f"(px[r]), "+f"(py[r]), "+f"(pz[r]));
            map_evalR<EX, R>(s_prog, fp.n_insn, fp.max_dist, px, py, pz, s, spill);

            if (!normal) {
                // ---------------- march update, branch-free per slot (wgsl:97-114) ----------------
                RM_FOR_R {
                    const bool live = (meta[r] & M_LIVE) != 0;
                    const bool hit = live && s[r] < fp.min_dist;  // wgsl:97
                    const unsigned hm = __ballot_sync(FULL, hit);
                    bool pushed = false;
                    if (hm != 0) {  // warp-uniform: hand the hitting rays to the normal queue
                        const int idx = hq_count + __popc(hm & lt_mask);
                        pushed = hit && idx < HQ_CAP;
                        if (pushed) ws.hitq[idx] = make_float4(px[r], py[r], pz[r], __uint_as_float(meta[r] & M_QB));
                        hq_count = min(hq_count + __popc(hm), HQ_CAP);
                    }
                    const bool far = s[r] > fp.max_dist;  // wgsl:109-111
                    const bool adv = live && !hit && !far;
                    const uint32_t m_adv = meta[r] + (1u << M_ITER_SHIFT);
                    const bool spent = (m_adv >> M_ITER_SHIFT) >= fp.max_iter;  // loop bound wgsl:90
                    if (adv) {
                        dist[r] = A::add(dist[r], s[r]);  // wgsl:114
                        meta[r] = m_adv;
                    }
                    const bool out = live && !hit && (far || spent);  // miss: floor / sky colour already parked
                    const uint32_t b = (meta[r] >> M_BUF_SHIFT) & 3u;
                    const uint32_t inc = (out ? 1u : 0u) << ((b & 1u) * 16u);
                    dec_lo += (b & 2u) ? 0u : inc;
                    dec_hi += (b & 2u) ? inc : 0u;
                    if (COUNT) {
                        if (live && (!hit || pushed)) ++c_step;  // a hit that found the queue full is re-evaluated
                        if (out && ws.codes[b][meta[r] & M_Q] != RM_CODE_SKY) ++c_floor;
                    }
                    if (pushed || out) meta[r] = 0;
                }
            } else {
                // ---------------- normal + diffuse term of the popped rays (wgsl:98-105,138-143) ----
#pragma unroll
                for (int j = 0; j < RPL; ++j) {
                    const int e = lane * RPL + j;
                    if (e < n_take) {
                        const float4 h = ws.hitq[hq_base + e];
                        const uint32_t bits = __float_as_uint(h.w);
                        const uint32_t b = (bits >> M_BUF_SHIFT) & 3u;
                        ws.codes[b][bits & M_Q] = wf_shade<EX>(s[4 * j], s[4 * j + 1], s[4 * j + 2], s[4 * j + 3], h.x, h.y, h.z);
                        const uint32_t inc = 1u << ((b & 1u) * 16u);
                        dec_lo += (b & 2u) ? 0u : inc;
                        dec_hi += (b & 2u) ? inc : 0u;
                        if (COUNT) ++c_hit;
                    }
                }
                hq_count = hq_base;
            }
            // fold the packed per-lane counts into the tiles' `remaining` (two warp reductions)
            const uint32_t dl = __reduce_add_sync(FULL, dec_lo), dh = __reduce_add_sync(FULL, dec_hi);
            if ((dl | dh) != 0) {  // warp-uniform
                if (lane < WF_K) {
                    const uint32_t w = (lane & 2) ? dh : dl;
                    ws.remaining[lane] -= (int)((w >> ((lane & 1) * 16)) & 0xffffu);
                }
                check = true;
            }
            __syncwarp();
        }
        // ---------------- resolve tiles whose 512 samples are all in ----------------
        if (check) {  // warp-uniform
            __syncwarp();
#pragma unroll 1
            for (int k = 0; k < WF_K; ++k) {
                const int tile = ws.tile[k];
                if (tile < 0 || ws.remaining[k] != 0) continue;  // warp-uniform
                const uint32_t pix = wf_resolve(ws.codes[k], lane);  // wgsl:44-45,68-75 in sample order
                const uint32_t tx = (uint32_t)tile % tiles_x, ty = (uint32_t)tile / tiles_x;
                const uint32_t x0 = tx * TILE_W, y0 = ty * TILE_H;
                if (x0 + TILE_W <= fp.W && y0 + TILE_H <= fp.nrows && (pitch & 15) == 0) {
                    // lanes 0..7 each gather four pixels of one half-row and issue one 128-bit store
                    const int row = (lane >> 1) & 3, half = lane & 1;
                    const int srcl = row * 8 + half * 4;
                    uint4 v;
                    v.x = __shfl_sync(FULL, pix, srcl);
                    v.y = __shfl_sync(FULL, pix, srcl + 1);
                    v.z = __shfl_sync(FULL, pix, srcl + 2);
                    v.w = __shfl_sync(FULL, pix, srcl + 3);
                    if (lane < 8)
                        *reinterpret_cast<uint4 *>(dst + (size_t)dst_row(fp, y0 + row) * pitch + (size_t)(x0 + half * 4) * 4) = v;
                } else {
                    const uint32_t x = x0 + (lane & 7), y = y0 + (lane >> 3);
                    if (x < fp.W && y < fp.nrows) *reinterpret_cast<uint32_t *>(dst + (size_t)dst_row(fp, y) * pitch + (size_t)x * 4) = pix;
                }
                __syncwarp();
                if (lane == 0) ws.tile[k] = -1;
            }
            __syncwarp();
        }
        // ---------------- termination: nothing marching, nothing queued, no tiles left ----------------
        if (!live_any && hq_count == 0 && ring_count == 0 && exhausted) {
            bool pending = false;
#pragma unroll
            for (int k = 0; k < WF_K; ++k) pending |= (ws.tile[k] >= 0);
            if (!pending) break;
        }
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

template <bool EX, int R, bool COUNT>
cudaError_t launch2_t(WavefrontState *ws, const FrameParams &fp, const RmInsn *prog, uint8_t *dst, size_t pitch,
                      unsigned long long *counters, unsigned int *tc, cudaStream_t st) {
    auto kern = k_wavefront2<EX, R, COUNT>;
    const size_t smem = (size_t)fp.n_insn * sizeof(RmInsn) + WF_WARPS * sizeof(WarpShared2<R>);
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int per_sm = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, WF_WARPS * 32, smem);
    if (e != cudaSuccess) return e;
    if (per_sm < 1) return cudaErrorLaunchOutOfResources;
    const uint32_t tiles = ((fp.W + TILE_W - 1) / TILE_W) * ((fp.nrows + TILE_H - 1) / TILE_H);
    unsigned grid = (unsigned)(ws->sm_count * per_sm);
    const unsigned need = (tiles + WF_WARPS - 1) / WF_WARPS;
    if (grid > need) grid = need;
    if (grid == 0) return cudaSuccess;
    kern<<<grid, WF_WARPS * 32, smem, st>>>(fp, prog, dst, pitch, tc, counters);
    return cudaGetLastError();
}

}  // namespace

cudaError_t launch_wavefront2(WavefrontState *ws, bool exact, const FrameParams &fp, const RmInsn *prog_dev,
                              uint8_t *dst, size_t pitch, unsigned long long *counters, cudaStream_t st) {
    if (fp.nrows == 0 || fp.W == 0) return cudaSuccess;
    unsigned int *tc = ws->d_tile_counters + (ws->ring_pos++ % RING);
    cudaError_t e = cudaMemsetAsync(tc, 0, sizeof(unsigned int), st);
    if (e != cudaSuccess) return e;
    static const int kR = [] {
        const char *s = std::getenv("RM_WF_R");
        const int v = s ? std::atoi(s) : 8;
        return (v == 4 || v == 8) ? v : 8;
    }();
#define RM_WF2_DISPATCH(EXV, RV)                                                                       \
    (counters ? launch2_t<EXV, RV, true>(ws, fp, prog_dev, dst, pitch, counters, tc, st)               \
              : launch2_t<EXV, RV, false>(ws, fp, prog_dev, dst, pitch, counters, tc, st))
    if (exact) return kR == 4 ? RM_WF2_DISPATCH(true, 4) : RM_WF2_DISPATCH(true, 8);
    return kR == 4 ? RM_WF2_DISPATCH(false, 4) : RM_WF2_DISPATCH(false, 8);
#undef RM_WF2_DISPATCH
}

}  // namespace rm
