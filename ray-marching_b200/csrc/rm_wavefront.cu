// rm_wavefront.cu — the production kernel of the render pass: persistent warps over 8x4 pixel
// tiles, each lane owning R ray slots that run a per-ray state machine
//     MARCH -> (hit) TAP0 -> TAP1 -> TAP2 -> TAP3 -> shade -> done        (wgsl:87-144)
// so that the only expensive thing — one map_scene evaluation of the decoded program per slot —
// is executed by all 32*R slots in lock-step, with warp-uniform control flow through the
// interpreter.  Finished slots are refilled from the warp's tile pool by ballot/popc ranking
// (compaction of live rays); a tile's 512 sample results are parked in shared memory and resolved
// in the shader's sample order, so the image is bit-identical to the one-thread-per-sample path.
// The framebuffer tile is written with 128-bit stores.
#include "rm_wavefront.h"
#include "rm_wavefront_common.cuh"

#include <cstdlib>

namespace rm {

namespace {

template <int R>
struct WarpShared {
    float codes[WF_K][TILE_RAYS];  // per-sample result codes; pre-set to the ray's miss code
    float4 ring[RING_N];           // generated rays waiting for a slot: dir.xyz + meta bits
    float taps[3][32 * R];         // first three normal taps of rays that hit (rarely touched)
    int remaining[WF_K];
    int tile[WF_K];  // tile id, or -1 when the buffer is free
};

template <bool EX, int R, bool COUNT>
__global__ void __launch_bounds__(WF_WARPS * 32, (R >= 8 ? 4 : (R == 4 ? 5 : 6)))
k_wavefront(const __grid_constant__ FrameParams fp, const RmInsn *__restrict__ prog_g, uint8_t *__restrict__ dst,
            size_t pitch, unsigned int *__restrict__ tile_counter, unsigned long long *__restrict__ counters) {
    using A = Arith<EX>;
    extern __shared__ float4 smem_f4[];
    float4 *s_prog = smem_f4;
    const uint32_t nq = fp.n_insn * 3;
    WarpShared<R> *s_warps = reinterpret_cast<WarpShared<R> *>(smem_f4 + nq);
    for (uint32_t i = threadIdx.x; i < nq; i += blockDim.x) s_prog[i] = reinterpret_cast<const float4 *>(prog_g)[i];
    const int lane = threadIdx.x & 31;
    WarpShared<R> &ws = s_warps[threadIdx.x >> 5];
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

    int cur_buf = 0, pool_next = TILE_RAYS, ring_head = 0, ring_count = 0;
    uint32_t cur_tile = 0;
    bool exhausted = false;
    uint32_t c_step = 0, c_hit = 0, c_floor = 0, c_sample = 0;

    for (;;) {
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
                } else {
                    if (COUNT && code != RM_CODE_SKY) ++c_floor;
                    atomicSub(&ws.remaining[cur_buf], 1);  // nothing to march
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
        // ---------------- termination ----------------
        bool live_any = false;
        RM_FOR_R live_any |= (meta[r] & M_LIVE) != 0;
        if (!__any_sync(FULL, live_any)) {
            __syncwarp();
            bool pending = false;
#pragma unroll
            for (int k = 0; k < WF_K; ++k) pending |= (ws.tile[k] >= 0);
            if (!pending && exhausted) break;
        }

        // ---------------- evaluation points ----------------
        float px[R], py[R], pz[R], s[R];
        RM_FOR_R {
            const uint32_t st = (meta[r] >> M_STATE_SHIFT) & 7u;
            const float bx = A::mad(rdx[r], dist[r], ro[0]), by = A::mad(rdy[r], dist[r], ro[1]),
                        bz = A::mad(rdz[r], dist[r], ro[2]);  // wgsl:91
            // tetrahedron taps wgsl:138-142: (+,-,-) (-,-,+) (-,+,-) (+,+,+)
            const float e = RM_NORMAL_EPS;
            const float ox = (st == 1 || st == 4) ? e : -e;
            const float oy = (st >= 3) ? e : -e;
            const float oz = (st == 2 || st == 4) ? e : -e;
            px[r] = st == 0 ? bx : __fadd_rn(bx, ox);
            py[r] = st == 0 ? by : __fadd_rn(by, oy);
            pz[r] = st == 0 ? bz : __fadd_rn(bz, oz);
            // keep the points in registers: without this ptxas rematerialises the select/add
            // chain above inside every interpreted instruction
            asm volatile("" : "+f"(px[r]), "+f"(py[r]), "+f"(pz[r]));
        }
        map_evalR<EX, R>(s_prog, fp.n_insn, fp.max_dist, px, py, pz, s, spill);

        // ---------------- per-slot state update ----------------
        RM_FOR_R {
            if (!(meta[r] & M_LIVE)) continue;
            const uint32_t st = (meta[r] >> M_STATE_SHIFT) & 7u;
            float code = 0.0f;
            bool done = false;
            if (st == 0) {
                if (COUNT) ++c_step;
                if (s[r] < fp.min_dist) {  // wgsl:97 -> start the normal taps
                    meta[r] += 1u << M_STATE_SHIFT;
                } else {
                    bool miss = s[r] > fp.max_dist;  // wgsl:109-111
                    if (!miss) {
                        dist[r] = A::add(dist[r], s[r]);  // wgsl:114
                        meta[r] += 1u << M_ITER_SHIFT;
                        miss = (meta[r] >> M_ITER_SHIFT) >= fp.max_iter;
                    }
                    if (miss) {  // floor / sky colour was parked at generation time
                        const uint32_t b = (meta[r] >> M_BUF_SHIFT) & 3u;
                        if (COUNT && ws.codes[b][meta[r] & M_Q] != RM_CODE_SKY) ++c_floor;
                        atomicSub(&ws.remaining[b], 1);
                        meta[r] = 0;
                    }
                }
            } else if (st < 4) {
                ws.taps[st - 1][r * 32 + lane] = s[r];
                meta[r] += 1u << M_STATE_SHIFT;
            } else {
                const float hx = A::mad(rdx[r], dist[r], ro[0]), hy = A::mad(rdy[r], dist[r], ro[1]),
                            hz = A::mad(rdz[r], dist[r], ro[2]);
                code = wf_shade<EX>(ws.taps[0][r * 32 + lane], ws.taps[1][r * 32 + lane], ws.taps[2][r * 32 + lane], s[r],
                                    hx, hy, hz);
                if (COUNT) ++c_hit;
                done = true;
            }
            if (done) {
                const uint32_t b = (meta[r] >> M_BUF_SHIFT) & 3u;
                ws.codes[b][meta[r] & M_Q] = code;
                atomicSub(&ws.remaining[b], 1);
                meta[r] = 0;
            }
        }
        __syncwarp();

        // ---------------- resolve tiles whose 512 samples are all in ----------------
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
cudaError_t launch_t(WavefrontState *ws, const FrameParams &fp, const RmInsn *prog, uint8_t *dst, size_t pitch,
                     unsigned long long *counters, unsigned int *tc, cudaStream_t st) {
    auto kern = k_wavefront<EX, R, COUNT>;
    const size_t smem = (size_t)fp.n_insn * sizeof(RmInsn) + WF_WARPS * sizeof(WarpShared<R>);
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

cudaError_t wavefront_init(WavefrontState *ws, int device) {
    ws->device = device;
    cudaError_t e = cudaDeviceGetAttribute(&ws->sm_count, cudaDevAttrMultiProcessorCount, device);
    if (e != cudaSuccess) return e;
    e = cudaMalloc(&ws->d_tile_counters, RING * sizeof(unsigned int));
    if (e != cudaSuccess) return e;
    return cudaMemset(ws->d_tile_counters, 0, RING * sizeof(unsigned int));
}

void wavefront_destroy(WavefrontState *ws) {
    if (ws->d_tile_counters) cudaFree(ws->d_tile_counters);
    ws->d_tile_counters = nullptr;
}

cudaError_t launch_wavefront(WavefrontState *ws, bool exact, const FrameParams &fp, const RmInsn *prog_dev,
                             uint8_t *dst, size_t pitch, unsigned long long *counters, cudaStream_t st) {
    if (fp.nrows == 0 || fp.W == 0) return cudaSuccess;
    unsigned int *tc = ws->d_tile_counters + (ws->ring_pos++ % RING);
    cudaError_t e = cudaMemsetAsync(tc, 0, sizeof(unsigned int), st);
    if (e != cudaSuccess) return e;
    static const int kR = [] {
        const char *s = std::getenv("RM_WF_R");
        const int v = s ? std::atoi(s) : 8;
        return (v == 1 || v == 2 || v == 4 || v == 8) ? v : 8;
    }();
#define RM_WF_DISPATCH(EXV, RV)                                                                        \
    (counters ? launch_t<EXV, RV, true>(ws, fp, prog_dev, dst, pitch, counters, tc, st)                \
              : launch_t<EXV, RV, false>(ws, fp, prog_dev, dst, pitch, counters, tc, st))
    if (exact) {
        if (kR == 1) return RM_WF_DISPATCH(true, 1);
        if (kR == 2) return RM_WF_DISPATCH(true, 2);
        if (kR == 8) return RM_WF_DISPATCH(true, 8);
        return RM_WF_DISPATCH(true, 4);
    } else {
        if (kR == 1) return RM_WF_DISPATCH(false, 1);
        if (kR == 2) return RM_WF_DISPATCH(false, 2);
        if (kR == 8) return RM_WF_DISPATCH(false, 8);
        return RM_WF_DISPATCH(false, 4);
    }
#undef RM_WF_DISPATCH
}

}  // namespace rm
