// rm_wavefront.h — launch wrapper of the persistent wavefront kernel (rm_wavefront.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "rm_device.cuh"

namespace rm {

struct WavefrontState {
    unsigned int *d_tile_counters = nullptr;  // ring of global tile-queue heads, one per launch in flight
    unsigned int ring_pos = 0;
    int sm_count = 0;
    int device = 0;
};

cudaError_t wavefront_init(WavefrontState *ws, int device);
void wavefront_destroy(WavefrontState *ws);

// Shades the band described by fp into dst (RGBA8, `pitch` bytes between rows).
// counters: optional device array {N_step, N_hit, N_floor, N_sample} (adds bookkeeping).
cudaError_t launch_wavefront(WavefrontState *ws, bool exact, const FrameParams &fp, const RmInsn *prog_dev,
                             uint8_t *dst, size_t pitch, unsigned long long *counters, cudaStream_t st);

// Same contract, second-generation kernel (rm_wavefront2.cu): march-only slots + hit queue.
cudaError_t launch_wavefront2(WavefrontState *ws, bool exact, const FrameParams &fp, const RmInsn *prog_dev,
                              uint8_t *dst, size_t pitch, unsigned long long *counters, cudaStream_t st);

// Shallow-program kernel (rm_direct.cu): 4 sample rays per thread, fused in-order resolve, no queues.
// early_out: opt-in exact early-out of the interpreter (rm_wavefront_common.cuh), bit-identical output.
// extended: the program holds N4 instructions (Program::extended) -> EXT interpreter (early_out is then ignored).
cudaError_t launch_direct(bool exact, bool early_out, bool extended, const FrameParams &fp, const RmInsn *prog_dev, uint8_t *dst,
                          size_t pitch, unsigned long long *counters, cudaStream_t st);

// Culled paint (rm_cull.cu, opt-in): the same kernel, every block running only the instructions its tile's mask keeps.
// direct_rays_for: the rays per thread launch_direct picks for this program (the mask tiles are 8 x direct_tile_rows(R)).
int direct_rays_for(const FrameParams &fp);
cudaError_t launch_direct_culled(bool exact, const FrameParams &fp, const RmInsn *prog_dev, uint8_t *dst, size_t pitch,
                                 unsigned long long *counters, const unsigned long long *masks, cudaStream_t st);

// The same kernel with the per-sample classes / distances of its own march written out (verification, rm_paint_aov).
cudaError_t launch_direct_aov(bool exact, bool extended, const FrameParams &fp, const RmInsn *prog_dev, uint8_t *dst, size_t pitch,
                              uint8_t *cls, float *t, cudaStream_t st);

}  // namespace rm
