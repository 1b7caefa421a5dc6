// hc_grid.h — TEST INFRASTRUCTURE ONLY: runs a kernel body on the host, one thread per lane, one warp (32 lanes in
// lock-step at the warp collectives of shim/cuda_runtime.h) at a time.  Blocks and warps are independent in the kernel
// bodies under test (no block-wide barrier, no shared memory written by the body), so warps are simply run one after
// the other, a few of them concurrently.
#pragma once
#include "cuda_runtime.h"  // the shim

#include <thread>

inline thread_local dim3 threadIdx, blockIdx, blockDim;
inline thread_local HcWarp *hc_warp = nullptr;
inline thread_local int hc_lane = 0;

inline thread_local dim3 gridDim;

template <class F>
void hc_run_grid(dim3 grid3, unsigned block_threads, F body) {
    const unsigned grid = grid3.x * grid3.y;
    const unsigned warps_per_block = block_threads / 32, n_warps = grid * warps_per_block;
    unsigned workers = std::thread::hardware_concurrency() / 2;
    workers = workers < 1 ? 1 : (workers > 8 ? 8 : workers);
    std::vector<std::thread> pool;
    for (unsigned wk = 0; wk < workers; ++wk) {
        pool.emplace_back([=, &body] {
            for (unsigned g = wk; g < n_warps; g += workers) {
                HcWarp warp;
                pthread_barrier_init(&warp.bar, nullptr, 32);
                std::thread lanes[32];
                for (int lane = 0; lane < 32; ++lane) {
                    lanes[lane] = std::thread([&, lane] {
                        threadIdx.x = (g % warps_per_block) * 32 + lane;
                        blockIdx.x = (g / warps_per_block) % grid3.x;
                        blockIdx.y = (g / warps_per_block) / grid3.x;
                        gridDim = grid3;
                        blockDim.x = block_threads;
                        hc_warp = &warp;
                        hc_lane = lane;
                        body();
                    });
                }
                for (auto &t : lanes) t.join();
                pthread_barrier_destroy(&warp.bar);
            }
        });
    }
    for (auto &t : pool) t.join();
}
