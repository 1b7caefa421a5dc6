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
    // fast mode: the affine form of wgsl:52-61.  p.x = (2/W) X - 1 with X = px + 0.5 + (i - 1.5) * W / (4 extent.x);
    // p.y = 1 - (2/H) Y with Y = py + 0.5 - (j - 1.5) * H / (4 extent.y); pt_world - ro = M (p.x, p.y, -1, 1) - ro with M = inv_view * inv_proj (column-major).
    {
        double M[16];
        for (int c = 0; c < 4; ++c)
            for (int r = 0; r < 4; ++r) {
                double a = 0.0;
                for (int k = 0; k < 4; ++k) a += (double)fp.inv_view[4 * k + r] * (double)fp.inv_proj[4 * c + k];
                M[4 * c + r] = a;
            }
        const double w2 = 2.0 / (double)W, h2 = 2.0 / (double)H;
        for (int r = 0; r < 4; ++r) {
            fp.ray_cx[r] = (float)(M[r] * w2);
            fp.ray_cy[r] = (float)(-M[4 + r] * h2);
            fp.ray_c0[r] = (float)(-M[r] + M[4 + r] - M[8 + r] + M[12 + r] - (double)fp.ro4[r]);
        }
        const double gx = (double)W / (4.0 * (double)fp.extent[0]), gy = (double)H / (4.0 * (double)fp.extent[1]);
        // a larger j moves the sample UP (larger p.y), i.e. towards smaller Y
        fp.ray_gx = (float)gx; fp.ray_gy = (float)-gy;
        fp.ray_bx = (float)(0.5 - 1.5 * gx); fp.ray_by = (float)(0.5 + 1.5 * gy);
    }
    return fp;
}

}  // namespace rm
