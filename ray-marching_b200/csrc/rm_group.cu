// rm_group.cu — the single-process multi-GPU surface of the render pass (SURVEY.md 8(b) "Multi-GPU (new surface)",
// 8(e)): one host thread — e.g. the reference's eframe main thread (main.rs:44-82) — drives every GPU of the box.
// A group is one rm_ctx per device plus a frame on member 0's device that the other members can address (peer access
// over NVLink).  It is written entirely on top of the public C ABI (rm_b200.h): the per-process form used under
// torchrun (rm_paint_strips_into + rm_shared_frame_*, rm_paint_poses_host) and this one share every kernel and store path.
//   config 4  rm_group_paint_bands: member i renders strips i, i+n, ... of ONE frame and its render kernel stores them
//             straight into member 0's frame at their frame rows — no gather, no re-order pass.
//   config 5  rm_group_paint_poses: pose p goes to member p % n; each member streams its frames to the caller's host
//             buffer (device ring + copy stream), members run concurrently on their own host threads.
#include <cuda_runtime.h>

#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../include/rm_b200.h"

struct rm_group {
    std::vector<rm_ctx *> ctx;
    std::vector<int> dev;
    uint8_t *frame = nullptr;  // on dev[0]
    size_t frame_cap = 0, frame_bytes = 0, frame_pitch = 0;
    std::string err;
};

namespace {

int gfail(rm_group *g, int code, const std::string &what) {
    if (g) g->err = what;
    return code;
}
int member_fail(rm_group *g, int i, int rc) {
    return gfail(g, rc, "member " + std::to_string(i) + " (device " + std::to_string(g->dev[i]) + "): " +
                            (rm_last_error(g->ctx[i]) ? rm_last_error(g->ctx[i]) : rm_strerror(rc)));
}

}  // namespace

extern "C" {

int rm_group_create(const int *cuda_devices, int n, rm_group **out) {
    if (!cuda_devices || n <= 0 || !out) return RM_E_INVALID_ARG;
    rm_group *g = new rm_group();
    for (int i = 0; i < n; ++i) {
        rm_ctx *c = nullptr;
        const int rc = rm_create(cuda_devices[i], &c);
        if (rc != RM_OK) {
            for (rm_ctx *p : g->ctx) rm_destroy(p);
            delete g;
            return rc;
        }
        g->ctx.push_back(c);
        g->dev.push_back(cuda_devices[i]);
    }
    // members on other devices store into member 0's frame: they need peer access to device 0
    for (int i = 1; i < n; ++i) {
        if (g->dev[i] == g->dev[0]) continue;
        int can = 0;
        cudaError_t e = cudaDeviceCanAccessPeer(&can, g->dev[i], g->dev[0]);
        if (e == cudaSuccess && can) {
            cudaSetDevice(g->dev[i]);
            e = cudaDeviceEnablePeerAccess(g->dev[0], 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); e = cudaSuccess; }
        }
        if (e != cudaSuccess || !can) {
            for (rm_ctx *p : g->ctx) rm_destroy(p);
            delete g;
            return RM_E_CUDA;
        }
    }
    *out = g;
    return RM_OK;
}

void rm_group_destroy(rm_group *g) {
    if (!g) return;
    for (rm_ctx *p : g->ctx) rm_sync(p);
    if (g->frame) { cudaSetDevice(g->dev[0]); cudaFree(g->frame); }
    for (rm_ctx *p : g->ctx) rm_destroy(p);
    delete g;
}

int rm_group_size(rm_group *g) { return g ? (int)g->ctx.size() : 0; }
rm_ctx *rm_group_ctx(rm_group *g, int i) { return (g && i >= 0 && i < (int)g->ctx.size()) ? g->ctx[i] : nullptr; }
const char *rm_group_last_error(rm_group *g) { return g ? g->err.c_str() : "null group"; }

int rm_group_set_mode(rm_group *g, int mode) {
    if (!g) return RM_E_INVALID_ARG;
    for (size_t i = 0; i < g->ctx.size(); ++i) { const int rc = rm_set_mode(g->ctx[i], mode); if (rc != RM_OK) return member_fail(g, (int)i, rc); }
    return RM_OK;
}
int rm_group_set_limits(rm_group *g, const void *limits12) {
    if (!g) return RM_E_INVALID_ARG;
    for (size_t i = 0; i < g->ctx.size(); ++i) { const int rc = rm_set_limits(g->ctx[i], limits12); if (rc != RM_OK) return member_fail(g, (int)i, rc); }
    return RM_OK;
}

int rm_group_specialize(rm_group *g) {
    if (!g) return RM_E_INVALID_ARG;
    for (size_t i = 0; i < g->ctx.size(); ++i) { const int rc = rm_specialize(g->ctx[i]); if (rc != RM_OK) return member_fail(g, (int)i, rc); }
    return RM_OK;
}

// The <= 1.2 KB of inputs are replicated to every member; unchanged words are not re-uploaded.
int rm_group_prepare(rm_group *g, const void *uniforms144, uint32_t cmd_count, const uint32_t *words, uint32_t n_words) {
    if (!g) return RM_E_INVALID_ARG;
    for (size_t i = 0; i < g->ctx.size(); ++i) {
        const int rc = rm_prepare(g->ctx[i], uniforms144, cmd_count, words, n_words);
        if (rc != RM_OK) return member_fail(g, (int)i, rc);
    }
    return RM_OK;
}

int rm_group_paint_bands(rm_group *g, uint32_t W, uint32_t H, uint32_t strip_rows, void *host_rgba8) {
    if (!g || W == 0 || H == 0 || strip_rows == 0) return RM_E_INVALID_ARG;
    const size_t pitch = ((size_t)W * 4 + 15) & ~(size_t)15, need = pitch * H;
    if (need > g->frame_cap) {
        cudaSetDevice(g->dev[0]);
        if (g->frame) cudaFree(g->frame);
        g->frame = nullptr; g->frame_cap = 0;
        if (cudaMalloc((void **)&g->frame, need) != cudaSuccess) return gfail(g, RM_E_CUDA, "cudaMalloc of the group frame failed");
        g->frame_cap = need;
    }
    g->frame_bytes = need; g->frame_pitch = pitch;
    const uint32_t n = (uint32_t)g->ctx.size();
    // launches are asynchronous: one host thread starts every member, then waits for all of them
    for (uint32_t i = 0; i < n; ++i) {
        const int rc = rm_paint_strips_into(g->ctx[i], W, H, strip_rows, i, n, g->frame, pitch, nullptr, nullptr);
        if (rc != RM_OK) return member_fail(g, (int)i, rc);
    }
    for (uint32_t i = 0; i < n; ++i) {
        const int rc = rm_sync(g->ctx[i]);
        if (rc != RM_OK) return member_fail(g, (int)i, rc);
    }
    if (host_rgba8) {
        cudaSetDevice(g->dev[0]);
        if (cudaMemcpy2D(host_rgba8, (size_t)W * 4, g->frame, pitch, (size_t)W * 4, H, cudaMemcpyDeviceToHost) != cudaSuccess)
            return gfail(g, RM_E_CUDA, "download of the group frame failed");
    }
    return RM_OK;
}

int rm_group_frame(rm_group *g, void **dev_ptr, size_t *pitch_bytes) {
    if (!g || !g->frame) return RM_E_INVALID_ARG;
    if (dev_ptr) *dev_ptr = g->frame;
    if (pitch_bytes) *pitch_bytes = g->frame_pitch;
    return RM_OK;
}

int rm_group_paint_poses(rm_group *g, const void *uniforms144_array, uint32_t n_poses, uint32_t W, uint32_t H,
                         void *host_rgba8, size_t host_frame_stride) {
    if (!g || !uniforms144_array || !host_rgba8 || host_frame_stride < (size_t)W * 4 * H) return RM_E_INVALID_ARG;
    const uint32_t n = (uint32_t)g->ctx.size();
    std::vector<std::vector<char>> u(n);
    const char *src = (const char *)uniforms144_array;
    for (uint32_t p = 0; p < n_poses; ++p) u[p % n].insert(u[p % n].end(), src + (size_t)p * RM_UNIFORMS_BYTES, src + (size_t)(p + 1) * RM_UNIFORMS_BYTES);
    std::vector<int> rc(n, RM_OK);
    std::vector<std::thread> th;
    for (uint32_t i = 0; i < n; ++i) {
        const uint32_t cnt = (uint32_t)(u[i].size() / RM_UNIFORMS_BYTES);
        if (cnt == 0) continue;
        // member i's k-th frame is pose i + k*n: host slot base + i*stride, then every n-th slot
        th.emplace_back([=, &u, &rc] {
            rc[i] = rm_paint_poses_host(g->ctx[i], u[i].data(), cnt, W, H, (uint8_t *)host_rgba8 + (size_t)i * host_frame_stride,
                                        host_frame_stride * n, 0);
        });
    }
    for (auto &t : th) t.join();
    for (uint32_t i = 0; i < n; ++i) if (rc[i] != RM_OK) return member_fail(g, (int)i, rc[i]);
    return RM_OK;
}

}  // extern "C"
