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

}  // namespace rm
