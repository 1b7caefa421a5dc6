// renderer.cpp — the render-pass host: Uniforms / RayMarchLimits byte layouts, RayMarchingResources
// and RayMarchingCallback::{prepare, paint} (src/ray_marching/renderer.rs), driving the C ABI of
// librm_b200.so instead of wgpu.
#include <cmath>
#include <cstring>
#include <stdexcept>
#include <string>

#include "../../include/rm_b200.h"
#include "rm_host.hpp"

namespace ray_marching {

std::array<uint8_t, 144> Uniforms::as_shader_bytes() const {
    // WGSL uniform layout: vec2<f32> @0 (align 8), mat4x4<f32> @16 (align 16), mat4x4<f32> @80
    std::array<uint8_t, 144> out{};
    std::memcpy(out.data() + 0, viewport_extent.data(), 8);
    std::memcpy(out.data() + 16, inv_proj.data(), 64);
    std::memcpy(out.data() + 80, inv_view.data(), 64);
    return out;
}

std::array<uint8_t, 12> RayMarchLimits::as_shader_bytes() const {
    std::array<uint8_t, 12> out{};
    std::memcpy(out.data() + 0, &min_dist, 4);
    std::memcpy(out.data() + 4, &max_dist, 4);
    std::memcpy(out.data() + 8, &max_iter, 4);
    return out;
}

static void check(rm_ctx *ctx, int rc, const char *what) {
    if (rc != RM_OK) throw std::runtime_error(std::string(what) + ": " + rm_strerror(rc) + " (" + (ctx ? rm_last_error(ctx) : "") + ")");
}

RayMarchingResources::RayMarchingResources(int cuda_device) {
    // renderer.rs:52-174: pipeline + the three buffers; limits get their constant values once
    check(nullptr, rm_create(cuda_device, &ctx_), "rm_create");
    const auto lim = RayMarchLimits{}.as_shader_bytes();  // renderer.rs:130-140
    check(ctx_, rm_set_limits(ctx_, lim.data()), "rm_set_limits");
}

RayMarchingResources::~RayMarchingResources() { rm_destroy(ctx_); }

RayMarchingCallback::RayMarchingCallback(float time, std::optional<csg::CSGNode> csg_node, std::array<float, 2> viewport,
                                         Camera camera)
    : time_(time), csg_node_(std::move(csg_node)), viewport_(viewport), camera_(camera) {}

Uniforms RayMarchingCallback::uniforms() const {
    Uniforms u;
    // renderer.rs:206-211
    const float frac_pi_4 = 0.78539816339744830962f;
    u.viewport_extent = viewport_;
    u.inv_proj = perspective_inverse(viewport_[0] / viewport_[1], frac_pi_4, 1.0f, 10000.0f);
    u.inv_view = mat4_inverse(camera_.view());
    return u;
}

void RayMarchingCallback::prepare(RayMarchingResources &resources) const {
    const auto bytes = uniforms().as_shader_bytes();  // renderer.rs:213-222
    csg::CSGCommandBufferBuilder builder;             // renderer.rs:224-227
    if (csg_node_) csg_node_->build_commands(builder);
    // renderer.rs:230-239: cmd_count at byte 0, words from byte 4
    check(resources.ctx(), rm_prepare(resources.ctx(), bytes.data(), builder.cmd_count, builder.buffer.data(),
                                      (uint32_t)builder.buffer.size()), "rm_prepare");
}

void RayMarchingCallback::paint(RayMarchingResources &resources, void *dst, size_t pitch, void *stream) const {
    const uint32_t W = (uint32_t)std::lround(viewport_[0]), H = (uint32_t)std::lround(viewport_[1]);
    check(resources.ctx(), rm_paint(resources.ctx(), W, H, 0, H, dst, pitch, stream), "rm_paint");  // renderer.rs:252-254
}

std::vector<uint8_t> RayMarchingCallback::render_to_host(RayMarchingResources &resources) const {
    prepare(resources);
    paint(resources);
    const uint32_t W = (uint32_t)std::lround(viewport_[0]), H = (uint32_t)std::lround(viewport_[1]);
    std::vector<uint8_t> out((size_t)W * H * 4);
    check(resources.ctx(), rm_download(resources.ctx(), out.data(), out.size()), "rm_download");
    return out;
}

}  // namespace ray_marching
