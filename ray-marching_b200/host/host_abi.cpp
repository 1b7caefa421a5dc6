// host_abi.cpp — extern "C" view of the C++ host (include/rm_host.h).
#include <stdexcept>
#include <cmath>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "../../include/rm_host.h"
#include "rm_host.hpp"

using namespace ray_marching;

namespace {
std::vector<csg::CSGNode> g_nodes;
std::vector<OrbitCameraController> g_ctls;
std::vector<CSGNodeGraph> g_graphs;
std::vector<std::unique_ptr<RayMarchingResources>> g_res;
std::string g_err;

bool node_ok(int h) { return h >= 0 && h < (int)g_nodes.size(); }
int add_node(csg::CSGNode n) {
    g_nodes.push_back(std::move(n));
    return (int)g_nodes.size() - 1;
}
csg::CSGNodeTemplate::Kind tkind(int k) { return static_cast<csg::CSGNodeTemplate::Kind>(k); }
Camera make_camera(const float p[3], const float q[4]) {
    Camera c;
    c.position = {p[0], p[1], p[2]};
    c.rotation = Quat{q[0], q[1], q[2], q[3]};
    return c;
}
template <typename F>
int guarded(F f) {
    try {
        return f();
    } catch (const std::exception &e) {
        g_err = e.what();
        return RMH_E_PANIC;
    }
}
}  // namespace

extern "C" {

void rmh_reset(void) {
    g_nodes.clear();
    g_ctls.clear();
    g_graphs.clear();
    g_res.clear();
    g_err.clear();
}
const char *rmh_last_error(void) { return g_err.c_str(); }

int rmh_node_sphere(const float c[3], float r) { return add_node(csg::CSGNode::Sphere({c[0], c[1], c[2]}, r)); }
int rmh_node_box(const float c[3], const float r[3]) { return add_node(csg::CSGNode::Box({c[0], c[1], c[2]}, {r[0], r[1], r[2]})); }
int rmh_node_binary(int kind, int a, int b, float k) {
    if (!node_ok(a) || !node_ok(b)) return RMH_E_INVALID;
    switch (kind) {
        case RMH_UNION: return add_node(csg::CSGNode::Union(g_nodes[a], g_nodes[b]));
        case RMH_SUBTRACTION: return add_node(csg::CSGNode::Subtraction(g_nodes[a], g_nodes[b]));
        case RMH_INTERSECTION: return add_node(csg::CSGNode::Intersection(g_nodes[a], g_nodes[b]));
        case RMH_SMOOTH_UNION: return add_node(csg::CSGNode::SmoothUnion(g_nodes[a], g_nodes[b], k));
        default: return RMH_E_INVALID;
    }
}
int rmh_node_plane(const float n[3], float offset) {
    return guarded([&] { return add_node(csg::CSGNode::Plane({n[0], n[1], n[2]}, offset)); });
}
int rmh_node_translation(int child, const float t[3]) {
    if (!node_ok(child)) return RMH_E_INVALID;
    return add_node(csg::CSGNode::Translation(g_nodes[child], {t[0], t[1], t[2]}));
}
int rmh_node_rotation(int child, const float q[4]) {
    if (!node_ok(child)) return RMH_E_INVALID;
    return add_node(csg::CSGNode::Rotation(g_nodes[child], Quat{q[0], q[1], q[2], q[3]}));
}
int rmh_node_rotation_matrix(int child, const float m[9]) {
    if (!node_ok(child)) return RMH_E_INVALID;
    std::array<float, 9> a;
    std::memcpy(a.data(), m, sizeof(float) * 9);
    return add_node(csg::CSGNode::RotationMatrix(g_nodes[child], a));
}
int rmh_node_scale(int child, float factor) {
    if (!node_ok(child)) return RMH_E_INVALID;
    return guarded([&] { return add_node(csg::CSGNode::Scale(g_nodes[child], factor)); });
}
int rmh_build_commands(int node, uint32_t *words, uint32_t cap, uint32_t *cmd_count) {
    if (!node_ok(node)) return RMH_E_INVALID;
    csg::CSGCommandBufferBuilder b;
    g_nodes[node].build_commands(b);
    if (cmd_count) *cmd_count = b.cmd_count;
    if (words) std::memcpy(words, b.buffer.data(), sizeof(uint32_t) * std::min<size_t>(cap, b.buffer.size()));
    return (int)b.buffer.size();
}

const char *rmh_template_name(int kind) { return csg::CSGNodeTemplate{tkind(kind)}.name(); }
int rmh_template_param_count(int kind) { return (int)csg::CSGNodeTemplate{tkind(kind)}.input_params().size(); }
const char *rmh_template_param(int kind, int i, int *tag, float value[3]) {
    static std::string name;
    auto ps = csg::CSGNodeTemplate{tkind(kind)}.input_params();
    if (i < 0 || i >= (int)ps.size()) return nullptr;
    name = ps[i].first;
    const ValueType &v = ps[i].second;
    if (tag) *tag = (int)v.tag;
    if (value) {
        value[0] = value[1] = value[2] = 0;
        if (v.tag == ValueType::Tag::Scalar) value[0] = v.scalar;
        if (v.tag == ValueType::Tag::Vec3) std::memcpy(value, v.vec3.data(), 12);
    }
    return name.c_str();
}
int rmh_template_evaluate(int kind, int n, const char *const *names, const int *tags, const float *values3, const int *nodes) {
    return guarded([&]() -> int {
        std::map<std::string, ValueType> p;
        for (int i = 0; i < n; ++i) {
            const float *v = values3 + 3 * i;
            if (tags[i] == RMH_VT_SCALAR) p[names[i]] = ValueType::Scalar(v[0]);
            else if (tags[i] == RMH_VT_VEC3) p[names[i]] = ValueType::Vec3v({v[0], v[1], v[2]});
            else p[names[i]] = ValueType::CSGNode(node_ok(nodes[i]) ? std::optional<csg::CSGNode>(g_nodes[nodes[i]]) : std::nullopt);
        }
        auto out = csg::CSGNodeTemplate{tkind(kind)}.evaluate(p);
        return out ? add_node(*out) : RMH_NONE;
    });
}

int rmh_orbit_new(const float t[3], float radius) {
    g_ctls.emplace_back(Vec3{t[0], t[1], t[2]}, radius);
    return (int)g_ctls.size() - 1;
}
#define CTL_OK(h) ((h) >= 0 && (h) < (int)g_ctls.size())
int rmh_orbit_update(int ctl, int event, float x, float y) {
    if (!CTL_OK(ctl) || event < 0 || event > 2) return RMH_E_INVALID;
    g_ctls[ctl].update(OrbitCameraControllerEvent{static_cast<OrbitCameraControllerEvent::Tag>(event), x, y});
    return 0;
}
int rmh_orbit_state(int ctl, float out[7]) {
    if (!CTL_OK(ctl)) return RMH_E_INVALID;
    const auto &c = g_ctls[ctl];
    out[0] = c.target[0]; out[1] = c.target[1]; out[2] = c.target[2];
    out[3] = c.pitch; out[4] = c.yaw; out[5] = c.radius; out[6] = 0;
    return 0;
}
int rmh_orbit_set(int ctl, float yaw, float pitch) {
    if (!CTL_OK(ctl)) return RMH_E_INVALID;
    g_ctls[ctl].yaw = yaw;
    g_ctls[ctl].pitch = pitch;
    return 0;
}
int rmh_orbit_camera(int ctl, float position[3], float q[4]) {
    if (!CTL_OK(ctl)) return RMH_E_INVALID;
    const Camera c = g_ctls[ctl].camera();
    std::memcpy(position, c.position.data(), 12);
    q[0] = c.rotation.w; q[1] = c.rotation.i; q[2] = c.rotation.j; q[3] = c.rotation.k;
    return 0;
}
int rmh_camera_view(const float p[3], const float q[4], float out16[16]) {
    const Mat4 m = make_camera(p, q).view();
    std::memcpy(out16, m.data(), 64);
    return 0;
}
int rmh_uniforms(const float p[3], const float q[4], float vw, float vh, uint8_t out144[144]) {
    RayMarchingCallback cb(0.0f, std::nullopt, {vw, vh}, make_camera(p, q));
    const auto b = cb.uniforms().as_shader_bytes();
    std::memcpy(out144, b.data(), 144);
    return 0;
}
int rmh_limits_default(uint8_t out12[12]) {
    const auto b = RayMarchLimits{}.as_shader_bytes();
    std::memcpy(out12, b.data(), 12);
    return 0;
}

#define GRAPH_OK(h) ((h) >= 0 && (h) < (int)g_graphs.size())
int rmh_graph_new(void) {
    g_graphs.emplace_back();
    return (int)g_graphs.size() - 1;
}
int rmh_graph_add_root(int g) { return GRAPH_OK(g) ? g_graphs[g].add_root() : RMH_E_INVALID; }
int rmh_graph_add_node(int g, int kind) {
    if (!GRAPH_OK(g) || kind < 0 || kind > RMH_SCALE) return RMH_E_INVALID;
    return g_graphs[g].add_node(csg::CSGNodeTemplate{tkind(kind)});
}
int rmh_graph_remove_node(int g, int node) {
    if (!GRAPH_OK(g)) return RMH_E_INVALID;
    return guarded([&] { g_graphs[g].remove_node(node); return 0; });
}
int rmh_graph_set_scalar(int g, int node, const char *param, float v) {
    if (!GRAPH_OK(g)) return RMH_E_INVALID;
    return guarded([&] { g_graphs[g].set_input(node, param, ValueType::Scalar(v)); return 0; });
}
int rmh_graph_set_vec3(int g, int node, const char *param, const float v[3]) {
    if (!GRAPH_OK(g)) return RMH_E_INVALID;
    return guarded([&] { g_graphs[g].set_input(node, param, ValueType::Vec3v({v[0], v[1], v[2]})); return 0; });
}
int rmh_graph_connect(int g, int from, int to, const char *param) {
    if (!GRAPH_OK(g)) return RMH_E_INVALID;
    try {
        g_graphs[g].connect(from, to, param);
    } catch (const std::invalid_argument &e) {  // DataType mismatch: the editor refuses the connection, nothing panics
        g_err = e.what();
        return RMH_E_INVALID;
    } catch (const std::exception &e) {
        g_err = e.what();
        return RMH_E_PANIC;
    }
    return 0;
}
int rmh_graph_disconnect(int g, int to, const char *param) {
    if (!GRAPH_OK(g)) return RMH_E_INVALID;
    return guarded([&] { g_graphs[g].disconnect(to, param); return 0; });
}
int rmh_graph_evaluate_root(int g) {
    if (!GRAPH_OK(g)) return RMH_E_INVALID;
    return guarded([&]() -> int {
        auto n = g_graphs[g].evaluate_root();
        return n ? add_node(*n) : RMH_NONE;
    });
}
int rmh_graph_last_evaluations(int g) { return GRAPH_OK(g) ? (int)g_graphs[g].evaluations_last() : RMH_E_INVALID; }

int rmh_resources_new(int cuda_device) {
    try {
        g_res.push_back(std::make_unique<RayMarchingResources>(cuda_device));
        return (int)g_res.size() - 1;
    } catch (const std::exception &e) {
        g_err = e.what();
        return RMH_E_DEVICE;
    }
}
void *rmh_resources_ctx(int res) { return (res >= 0 && res < (int)g_res.size()) ? g_res[res]->ctx() : nullptr; }
int rmh_callback_render(int res, int node, const float p[3], const float q[4], float vw, float vh, uint8_t *out) {
    if (res < 0 || res >= (int)g_res.size()) return RMH_E_INVALID;
    return guarded([&]() -> int {
        RayMarchingCallback cb(0.0f, node_ok(node) ? std::optional<csg::CSGNode>(g_nodes[node]) : std::nullopt, {vw, vh},
                               make_camera(p, q));
        if (out) {
            const auto img = cb.render_to_host(*g_res[res]);
            std::memcpy(out, img.data(), img.size());
        } else {
            cb.prepare(*g_res[res]);
            cb.paint(*g_res[res]);
        }
        return 0;
    });
}

}  // extern "C"
