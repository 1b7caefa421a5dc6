// rm_kernels.h — launch wrappers of the sm_100a kernels (implemented in rm_kernels.cu and
// rm_wavefront.cu).  Host-callable, plain pointers; every wrapper returns a cudaError_t.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "rm_device.cuh"

namespace rm {

struct DebugOut {
    uint8_t *cls;                  // per sample class, or null
    float *t;                      // per sample distance, or null
    unsigned long long *counters;  // {N_step, N_hit, N_floor, N_sample}, or null
};

// Simple path: one thread per sample ray, then an in-order per-pixel resolve.
// extended: the program holds N4 instructions (Program::extended) -> EXT interpreter instantiation.
cudaError_t launch_samples_simple(bool exact, bool extended, const FrameParams &fp, const RmInsn *prog_dev, float *codes_dev,
                                  DebugOut dbg, cudaStream_t st);
cudaError_t launch_resolve(const FrameParams &fp, const float *codes_dev, uint8_t *dst, size_t pitch,
                           cudaStream_t st);
cudaError_t launch_map(bool exact, bool extended, const FrameParams &fp, const RmInsn *prog_dev, const float *pts_dev,
                       unsigned long long n, float *out_dev, cudaStream_t st);
// FFMA throughput probe: each thread runs `iters` x 8 independent FFMA chains.
cudaError_t launch_ffma_probe(int blocks, int threads, int iters, float *sink_dev, cudaStream_t st);

}  // namespace rm
