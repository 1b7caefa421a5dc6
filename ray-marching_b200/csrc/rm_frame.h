// rm_frame.h — host side: builds the per-launch FrameParams block (rm_device.cuh) from the reference's two uniform
// buffers (Uniforms 144 B, renderer.rs:29-34; RayMarchLimits 12 B, renderer.rs:36-41) and the band geometry.
// The per-frame constants of ray generation are computed here once, with the same IEEE float operations the shader's
// expressions denote (wgsl:39-40, 52-53), so that they are bit-identical to evaluating them per ray.  Compile the
// including file with -ffp-contract=off; volatile keeps the compiler from folding x * 0.
#pragma once
#include <cstring>

#include "rm_device.cuh"

namespace rm {

inline FrameParams make_frame_params(const float *uniforms36, float min_dist, float max_dist, uint32_t max_iter, uint32_t n_insn,
                                     uint32_t W, uint32_t H, uint32_t row0, uint32_t nrows) {
    FrameParams fp;
    std::memcpy(fp.extent, uniforms36, 8);
    std::memcpy(fp.inv_proj, uniforms36 + 4, 64);
    std::memcpy(fp.inv_view, uniforms36 + 20, 64);
    fp.min_dist = min_dist;
    fp.max_dist = max_dist;
    fp.max_iter = max_iter;
    fp.W = W; fp.H = H; fp.row0 = row0; fp.nrows = nrows;
    fp.strip_rows = 0; fp.strip_pitch = 0;
    fp.dst_frame_rows = 0;
    fp.n_insn = n_insn;
    // per-frame ray-generation constants: the shader's expressions, one IEEE float operation at a time
    for (int r = 0; r < 4; ++r) {
        volatile float acc = fp.inv_view[r] * 0.0f;
        acc = acc + fp.inv_view[4 + r] * 0.0f;
        acc = acc + fp.inv_view[8 + r] * 0.0f;
        acc = acc + fp.inv_view[12 + r] * 1.0f;
        fp.ro4[r] = acc;
    }
    for (int i = 0; i < 4; ++i) {
        volatile float off = ((float)i + 0.5f) / 4.0f - 0.5f;
        volatile float qx = off / fp.extent[0], qy = off / fp.extent[1];
        fp.sample_ax[i] = qx * 2.0f;
        fp.sample_ay[i] = qy * 2.0f;
    }
    return fp;
}

}  // namespace rm
