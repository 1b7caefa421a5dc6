// rm_host.hpp — C++ host mirroring the reference's Rust API surface for the render pass
// (Mesoptier/ray-marching; Rust is not available in this image, so the host above the C ABI is C++).
// Same names, argument meaning and error behaviour as the Rust items cited on each declaration,
// so that code and tests written against it read like the reference's own.
//
//   csg::CSGCommandType / CSGCommandBufferBuilder   src/ray_marching/csg/builder.rs:1-62
//   csg::CSGNode (+ Sphere, Box, Union, Subtraction) src/ray_marching/csg/mod.rs:28-45, primitives/*, operations/mod.rs
//   csg::CSGNodeTemplate, ValueType                  csg/mod.rs:21-26,47-65; csg_node_graph.rs:24-64
//   Camera, OrbitCameraController                    src/camera.rs:1-85
//   Uniforms, RayMarchLimits, RayMarchingResources,
//   RayMarchingCallback{prepare,paint}               src/ray_marching/renderer.rs:29-256
//   CSGNodeGraph::evaluate_root + Evaluator          src/csg_node_graph.rs:235-309   (headless: no egui)
// Extensions not in the reference: Intersection (opcode 102), SmoothUnion (103, k); and N4
// (SURVEY.md 8(f)): the node kinds the reference's enums reserve in comments (csg/mod.rs:34,41-44, builder.rs:8,16-23):
// Plane, Translation, Rotation, Scale.
#pragma once
#include <array>
#include <cstdint>
#include <map>
#include <memory>
#include <optional>
#include <string>
#include <utility>
#include <vector>

struct rm_ctx;

namespace ray_marching {

using Vec3 = std::array<float, 3>;
struct Quat { float w = 1, i = 0, j = 0, k = 0; };  // nalgebra UnitQuaternion<f32>

namespace csg {

// builder.rs:1-24 (#[repr(u32)])
enum class CSGCommandType : uint32_t {
    Sphere = 0, Box = 1, Plane = 2,                                           // builder.rs:6-8
    Union = 100, Subtraction = 101, Intersection = 102, SmoothUnion = 103,    // builder.rs:12-14
    TranslationPush = 200, TranslationPop = 201, RotationPush = 202, RotationPop = 203, ScalePush = 204, ScalePop = 205,  // :18-23
};

// builder.rs:26-62
struct CSGCommandBufferBuilder {
    uint32_t cmd_count = 0;
    std::vector<uint32_t> buffer;
    CSGCommandBufferBuilder &push_command(CSGCommandType t);   // builder.rs:41-45
    CSGCommandBufferBuilder &push_param_vec3(const Vec3 &v);   // builder.rs:49-54
    CSGCommandBufferBuilder &push_param_float(float v);        // builder.rs:58-61
};

struct CSGNode;
using NodeBox = std::shared_ptr<const CSGNode>;  // Box<CSGNode>; immutable, so sharing == cloning

// csg/mod.rs:28-45 (enum_dispatch over the node structs) + BuildCommands (csg/mod.rs:16-19)
struct CSGNode {
    enum class Kind { Sphere, Box, Union, Subtraction, Intersection, SmoothUnion, Plane, Translation, Rotation, Scale };
    Kind kind = Kind::Sphere;
    Vec3 center{0, 0, 0};   // sphere.rs:11, box.rs:10; Plane: unit normal; Translation: offset
    Vec3 radius{1, 1, 1};   // sphere: radius[0] (sphere.rs:12); box: radius xyz (box.rs:11)
    float k = 0.25f;        // SmoothUnion: k; Plane: offset h; Scale: factor
    std::array<float, 9> rotation{1, 0, 0, 0, 1, 0, 0, 0, 1};  // Rotation: row-major matrix, object -> world
    NodeBox a, b;           // binary operators: .0 and .1 (operations/mod.rs:9); space transforms: a = the one child

    static CSGNode Sphere(const Vec3 &center, float radius);
    static CSGNode Box(const Vec3 &center, const Vec3 &radius);
    static CSGNode Union(CSGNode a, CSGNode b);
    static CSGNode Subtraction(CSGNode a, CSGNode b);
    static CSGNode Intersection(CSGNode a, CSGNode b);
    static CSGNode SmoothUnion(CSGNode a, CSGNode b, float k);
    // N4.  Plane: dot(p, n) + offset with n normalised here (throws std::invalid_argument for a zero / non-finite
    // normal).  Space transforms have one child ("1 child, transforms space", builder.rs:16-17) and flatten to
    // Push, child, Pop.  Scale is uniform and must be finite and > 0 (throws otherwise).
    static CSGNode Plane(const Vec3 &normal, float offset);
    static CSGNode Translation(CSGNode child, const Vec3 &offset);
    static CSGNode Rotation(CSGNode child, const Quat &q);                       // unit quaternion, as camera.rs uses
    static CSGNode RotationMatrix(CSGNode child, const std::array<float, 9> &m);  // row-major, object -> world
    static CSGNode Scale(CSGNode child, float factor);
    // post-order walk: A, B, then the operator (operations/mod.rs:13-17); primitives: opcode then
    // params (sphere.rs:15-21, box.rs:14-20)
    void build_commands(CSGCommandBufferBuilder &builder) const;
    bool is_binary() const { return kind == Kind::Union || kind == Kind::Subtraction || kind == Kind::Intersection || kind == Kind::SmoothUnion; }
    bool is_transform() const { return kind == Kind::Translation || kind == Kind::Rotation || kind == Kind::Scale; }
};

}  // namespace csg

// csg_node_graph.rs:24-64
struct ValueType {
    enum class Tag { Scalar, Vec3, CSGNode };
    Tag tag = Tag::Scalar;
    float scalar = 0.0f;  // Default = Scalar(0.0), csg_node_graph.rs:66-70
    Vec3 vec3{0, 0, 0};
    std::optional<csg::CSGNode> node;
    static ValueType Scalar(float x);
    static ValueType Vec3v(const Vec3 &v);
    static ValueType CSGNode(std::optional<csg::CSGNode> n);
    std::optional<float> to_scalar() const;            // :44-49
    std::optional<Vec3> to_vec3() const;               // :51-56
    std::optional<csg::CSGNode> to_csg_node() const;   // :58-63
};

namespace csg {
// csg/mod.rs:21-26,47-65 — the editor's node templates (egui's InputParamKind is UI-only and dropped)
struct CSGNodeTemplate {
    enum class Kind { Sphere, Box, Union, Subtraction, Intersection, SmoothUnion, Plane, Translation, Rotation, Scale };
    Kind kind;
    const char *name() const;                                             // "Sphere", "Box", "Union", "Subtraction", ...
    std::vector<std::pair<std::string, ValueType>> input_params() const;  // names + defaults (sphere.rs:28-43 ...)
    // None when a connection-only input is missing (operations/mod.rs:47-48); throws std::runtime_error
    // where the reference unwrap()s (wrong value type / missing key: sphere.rs:47-48, box.rs:46-47)
    std::optional<CSGNode> evaluate(const std::map<std::string, ValueType> &input_params) const;
    static std::vector<CSGNodeTemplate> all();  // csg/mod.rs:56-64 (+ the two extension operators + the four N4 kinds)
};
}  // namespace csg

// ---- camera.rs -------------------------------------------------------------------------------
Quat quat_from_euler_angles(float roll, float pitch, float yaw);  // UnitQuaternion::from_euler_angles
Vec3 quat_rotate(const Quat &q, const Vec3 &v);
Quat quat_from_axis_angle(const Vec3 &axis, float angle);        // UnitQuaternion::from_axis_angle (axis normalised here)
std::array<float, 9> quat_to_matrix3(const Quat &q);             // UnitQuaternion::to_rotation_matrix, row-major
using Mat4 = std::array<float, 16>;  // column-major, like nalgebra storage / WGSL mat4x4

struct Camera {                // camera.rs:3-13
    Vec3 position{0, 0, 0};
    Quat rotation;
    Mat4 view() const;         // rotation.inverse() * Translation(-position), camera.rs:10-12
};

struct OrbitCameraControllerEvent {  // camera.rs:15-19
    enum class Tag { Pan, Orbit, Dolly } tag;
    float x = 0, y = 0;              // Pan([dx,dy]) / Orbit([dx,dy]) / Dolly(x)
};

struct OrbitCameraController {       // camera.rs:21-85
    Vec3 target;
    float pitch = 0, yaw = 0, radius;
    float pan_speed = 0.01f, yaw_speed = 0.01f, pitch_speed = 0.01f, dolly_speed = 0.01f;
    OrbitCameraController(const Vec3 &target, float radius);  // :38-50
    Quat rotation() const;                                    // :52-54
    Camera camera() const;                                    // :56-60
    void update(const OrbitCameraControllerEvent &e);         // :62-84
};

// ---- renderer.rs -----------------------------------------------------------------------------
Mat4 perspective_inverse(float aspect, float fovy, float znear, float zfar);  // Perspective3::new(..).inverse()
Mat4 mat4_inverse(const Mat4 &m);                                             // Affine3::inverse()

struct Uniforms {  // renderer.rs:29-34
    std::array<float, 2> viewport_extent{0, 0};
    Mat4 inv_proj{}, inv_view{};
    std::array<uint8_t, 144> as_shader_bytes() const;  // encase UniformBuffer layout, renderer.rs:21-27
};
struct RayMarchLimits {  // renderer.rs:36-41
    float min_dist = 0.01f, max_dist = 100.0f;
    uint32_t max_iter = 100;  // renderer.rs:133-137
    std::array<uint8_t, 12> as_shader_bytes() const;
};

// renderer.rs:43-175.  Owns the device-side resources (an rm_ctx of librm_b200.so).
// Throws std::runtime_error where the reference would panic (no device, CUDA failure).
class RayMarchingResources {
public:
    explicit RayMarchingResources(int cuda_device = 0);
    ~RayMarchingResources();
    RayMarchingResources(const RayMarchingResources &) = delete;
    RayMarchingResources &operator=(const RayMarchingResources &) = delete;
    rm_ctx *ctx() const { return ctx_; }
private:
    rm_ctx *ctx_ = nullptr;
};

// renderer.rs:177-256
class RayMarchingCallback {
public:
    RayMarchingCallback(float time, std::optional<csg::CSGNode> csg_node, std::array<float, 2> viewport, Camera camera);
    // renderer.rs:196-242: builds Uniforms + flattens the tree and uploads both
    void prepare(RayMarchingResources &resources) const;
    // renderer.rs:244-255: draws into `dst_rgba8_dev` (device; null = resources-owned buffer).
    // The viewport size in pixels is the callback rect (pixels_per_point = 1).
    void paint(RayMarchingResources &resources, void *dst_rgba8_dev = nullptr, size_t pitch_bytes = 0,
               void *cuda_stream = nullptr) const;
    // convenience: prepare + paint + read back, row-major RGBA8
    std::vector<uint8_t> render_to_host(RayMarchingResources &resources) const;
    Uniforms uniforms() const;  // what prepare() uploads
private:
    float time_;  // stored, never read (renderer.rs:178)
    std::optional<csg::CSGNode> csg_node_;
    std::array<float, 2> viewport_;
    Camera camera_;
};

// ---- csg_node_graph.rs (headless) -------------------------------------------------------------
// The editor graph as plain data: nodes instantiated from templates, inputs holding either a
// constant ValueType or a connection to another node's "SDF" output; a Root node with one "SDF" input.
class CSGNodeGraph {
public:
    using NodeId = int;
    NodeId add_root();                                   // NodeTemplate::Root, csg_node_graph.rs:130-139
    NodeId add_node(const csg::CSGNodeTemplate &t);      // build_node, :140-152 (inputs get template defaults)
    void remove_node(NodeId id);
    void set_input(NodeId node, const std::string &param, const ValueType &value);
    void connect(NodeId from_node, NodeId to_node, const std::string &to_param);  // output "SDF" -> input
    void disconnect(NodeId to_node, const std::string &to_param);
    std::optional<csg::CSGNode> evaluate_root() const;   // :251-261 (None: no Root / unconnected / missing input)
    size_t evaluations_last() const { return evals_; }   // template.evaluate calls of the last evaluate_root (memoised)
private:
    struct Input { std::string name; ValueType value; NodeId connection = -1; };
    struct Node { bool alive = false; bool root = false; csg::CSGNodeTemplate tmpl{csg::CSGNodeTemplate::Kind::Sphere}; std::vector<Input> inputs; };
    std::vector<Node> nodes_;
    mutable size_t evals_ = 0;
    friend struct Evaluator;
};

}  // namespace ray_marching
