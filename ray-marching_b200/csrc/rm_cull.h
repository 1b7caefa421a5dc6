// rm_cull.h — per-tile conservative culling of the decoded program (rm_cull.cu): opt-in, exact, never the headline.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "rm_device.cuh"

namespace rm {

// One 64-bit mask per pixel tile (8 wide, tile_h rows, the production kernel's block tile): bit i set = instruction i of
// the decoded program (fp.n_insn <= 64, reference instruction set) can change the accumulator somewhere a ray of the tile
// goes and is kept.  t_end: an upper bound of the ray parameter of any position a ray of the frame can visit.
cudaError_t launch_cull_masks(const FrameParams &fp, const RmInsn *prog_dev, uint32_t tile_h, float t_end, unsigned long long *masks_dev,
                              cudaStream_t st);

// Allocation-free variant (rm_paint_culled_static): the masks live in a file-scope static __device__ array of
// RM_CULL_STATIC_MAX_TILES entries (2 MiB per GPU that loads the library) instead of a cudaMalloc'ed buffer.  `device` must
// be the current device; returns cudaErrorInvalidValue when the band has more tiles than the array holds.
#ifndef RM_CULL_STATIC_MAX_TILES  /* the same value as include/rm_b200.h declares */
#define RM_CULL_STATIC_MAX_TILES 262144u
#endif
cudaError_t launch_cull_masks_static(int device, const FrameParams &fp, const RmInsn *prog_dev, uint32_t tile_h, float t_end,
                                     const unsigned long long **masks_out, cudaStream_t st);

}  // namespace rm
