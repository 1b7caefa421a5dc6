// csg.cpp — CSG node model, command-buffer builder and node templates (see rm_host.hpp for the
// reference lines each item mirrors).
#include <cmath>
#include <cstring>
#include <stdexcept>

#include "rm_host.hpp"

namespace ray_marching {

namespace csg {

static inline uint32_t to_bits(float f) {  // f32::to_bits, builder.rs:51,59
    uint32_t u;
    std::memcpy(&u, &f, 4);
    return u;
}

CSGCommandBufferBuilder &CSGCommandBufferBuilder::push_command(CSGCommandType t) {
    cmd_count += 1;
    buffer.push_back(static_cast<uint32_t>(t));
    return *this;
}
CSGCommandBufferBuilder &CSGCommandBufferBuilder::push_param_vec3(const Vec3 &v) {
    for (float x : v) buffer.push_back(to_bits(x));
    return *this;
}
CSGCommandBufferBuilder &CSGCommandBufferBuilder::push_param_float(float v) {
    buffer.push_back(to_bits(v));
    return *this;
}

CSGNode CSGNode::Sphere(const Vec3 &center, float radius) {
    CSGNode n;
    n.kind = Kind::Sphere;
    n.center = center;
    n.radius = {radius, radius, radius};
    return n;
}
CSGNode CSGNode::Box(const Vec3 &center, const Vec3 &radius) {
    CSGNode n;
    n.kind = Kind::Box;
    n.center = center;
    n.radius = radius;
    return n;
}
static CSGNode binary(CSGNode::Kind kind, CSGNode a, CSGNode b, float k = 0.25f) {
    CSGNode n;
    n.kind = kind;
    n.k = k;
    n.a = std::make_shared<const CSGNode>(std::move(a));
    n.b = std::make_shared<const CSGNode>(std::move(b));
    return n;
}
CSGNode CSGNode::Union(CSGNode a, CSGNode b) { return binary(Kind::Union, std::move(a), std::move(b)); }
CSGNode CSGNode::Subtraction(CSGNode a, CSGNode b) { return binary(Kind::Subtraction, std::move(a), std::move(b)); }
CSGNode CSGNode::Intersection(CSGNode a, CSGNode b) { return binary(Kind::Intersection, std::move(a), std::move(b)); }
CSGNode CSGNode::SmoothUnion(CSGNode a, CSGNode b, float k) { return binary(Kind::SmoothUnion, std::move(a), std::move(b), k); }

CSGNode CSGNode::Plane(const Vec3 &normal, float offset) {
    const float len = std::sqrt(normal[0] * normal[0] + normal[1] * normal[1] + normal[2] * normal[2]);
    if (!(len > 0.0f) || !std::isfinite(len)) throw std::invalid_argument("Plane: normal must be finite and non-zero");
    CSGNode n;
    n.kind = Kind::Plane;
    n.center = {normal[0] / len, normal[1] / len, normal[2] / len};
    n.k = offset;
    return n;
}
static CSGNode unary(CSGNode::Kind kind, CSGNode child) {
    CSGNode n;
    n.kind = kind;
    n.a = std::make_shared<const CSGNode>(std::move(child));
    return n;
}
CSGNode CSGNode::Translation(CSGNode child, const Vec3 &offset) {
    CSGNode n = unary(Kind::Translation, std::move(child));
    n.center = offset;
    return n;
}
CSGNode CSGNode::RotationMatrix(CSGNode child, const std::array<float, 9> &m) {
    CSGNode n = unary(Kind::Rotation, std::move(child));
    n.rotation = m;
    return n;
}
CSGNode CSGNode::Rotation(CSGNode child, const Quat &q) { return RotationMatrix(std::move(child), quat_to_matrix3(q)); }
CSGNode CSGNode::Scale(CSGNode child, float factor) {
    if (!(factor > 0.0f) || !std::isfinite(factor)) throw std::invalid_argument("Scale: factor must be finite and > 0");
    CSGNode n = unary(Kind::Scale, std::move(child));
    n.k = factor;
    return n;
}

void CSGNode::build_commands(CSGCommandBufferBuilder &builder) const {
    switch (kind) {
        case Kind::Plane:  // N4: opcode, then params (the order of sphere.rs:15-21)
            builder.push_command(CSGCommandType::Plane).push_param_vec3(center).push_param_float(k);
            return;
        case Kind::Translation:  // Push, the one child, Pop (builder.rs:16-19)
            builder.push_command(CSGCommandType::TranslationPush).push_param_vec3(center);
            a->build_commands(builder);
            builder.push_command(CSGCommandType::TranslationPop);
            return;
        case Kind::Rotation:  // the matrix row by row (builder.rs:20-21)
            builder.push_command(CSGCommandType::RotationPush)
                .push_param_vec3({rotation[0], rotation[1], rotation[2]})
                .push_param_vec3({rotation[3], rotation[4], rotation[5]})
                .push_param_vec3({rotation[6], rotation[7], rotation[8]});
            a->build_commands(builder);
            builder.push_command(CSGCommandType::RotationPop);
            return;
        case Kind::Scale:  // the Pop repeats the factor: the child's distance is multiplied by it (builder.rs:22-23)
            builder.push_command(CSGCommandType::ScalePush).push_param_float(k);
            a->build_commands(builder);
            builder.push_command(CSGCommandType::ScalePop).push_param_float(k);
            return;
        case Kind::Sphere:  // sphere.rs:15-21
            builder.push_command(CSGCommandType::Sphere).push_param_vec3(center).push_param_float(radius[0]);
            return;
        case Kind::Box:  // box.rs:14-20
            builder.push_command(CSGCommandType::Box).push_param_vec3(center).push_param_vec3(radius);
            return;
        default:
            break;
    }
    a->build_commands(builder);  // operations/mod.rs:14
    b->build_commands(builder);  // operations/mod.rs:15
    switch (kind) {              // operations/mod.rs:16
        case Kind::Union: builder.push_command(CSGCommandType::Union); break;
        case Kind::Subtraction: builder.push_command(CSGCommandType::Subtraction); break;
        case Kind::Intersection: builder.push_command(CSGCommandType::Intersection); break;
        default: builder.push_command(CSGCommandType::SmoothUnion).push_param_float(k); break;  // opcode, then param
    }
}

const char *CSGNodeTemplate::name() const {
    switch (kind) {
        case Kind::Sphere: return "Sphere";            // sphere.rs:26-28
        case Kind::Box: return "Box";                  // box.rs:25-27
        case Kind::Union: return "Union";              // stringify!($name), operations/mod.rs:25-27
        case Kind::Subtraction: return "Subtraction";
        case Kind::Intersection: return "Intersection";
        case Kind::Plane: return "Plane";
        case Kind::Translation: return "Translation";
        case Kind::Rotation: return "Rotation";
        case Kind::Scale: return "Scale";
        default: return "SmoothUnion";
    }
}

std::vector<std::pair<std::string, ValueType>> CSGNodeTemplate::input_params() const {
    switch (kind) {
        case Kind::Sphere:  // sphere.rs:30-43
            return {{"center", ValueType::Vec3v({0, 0, 0})}, {"radius", ValueType::Scalar(1.0f)}};
        case Kind::Box:  // box.rs:29-42
            return {{"center", ValueType::Vec3v({0, 0, 0})}, {"radius", ValueType::Vec3v({1, 1, 1})}};
        case Kind::SmoothUnion:
            return {{"A", ValueType::CSGNode(std::nullopt)}, {"B", ValueType::CSGNode(std::nullopt)}, {"k", ValueType::Scalar(0.25f)}};
        case Kind::Plane:
            return {{"normal", ValueType::Vec3v({0, 1, 0})}, {"offset", ValueType::Scalar(0.0f)}};
        case Kind::Translation:
            return {{"A", ValueType::CSGNode(std::nullopt)}, {"offset", ValueType::Vec3v({0, 0, 0})}};
        case Kind::Rotation:
            return {{"A", ValueType::CSGNode(std::nullopt)}, {"axis", ValueType::Vec3v({0, 1, 0})}, {"angle", ValueType::Scalar(0.0f)}};
        case Kind::Scale:
            return {{"A", ValueType::CSGNode(std::nullopt)}, {"factor", ValueType::Scalar(1.0f)}};
        default:  // operations/mod.rs:29-42
            return {{"A", ValueType::CSGNode(std::nullopt)}, {"B", ValueType::CSGNode(std::nullopt)}};
    }
}

static const ValueType &get(const std::map<std::string, ValueType> &m, const char *key) {
    auto it = m.find(key);
    if (it == m.end()) throw std::runtime_error(std::string("missing input param: ") + key);  // .get(..).unwrap()
    return it->second;
}
template <typename T>
static T unwrap(std::optional<T> o, const char *what) {
    if (!o) throw std::runtime_error(std::string("wrong value type for input param: ") + what);  // .unwrap()
    return *o;
}

std::optional<CSGNode> CSGNodeTemplate::evaluate(const std::map<std::string, ValueType> &p) const {
    switch (kind) {
        case Kind::Sphere:  // sphere.rs:46-50
            return CSGNode::Sphere(unwrap(get(p, "center").to_vec3(), "center"), unwrap(get(p, "radius").to_scalar(), "radius"));
        case Kind::Box:  // box.rs:45-49
            return CSGNode::Box(unwrap(get(p, "center").to_vec3(), "center"), unwrap(get(p, "radius").to_vec3(), "radius"));
        case Kind::Plane:
            return CSGNode::Plane(unwrap(get(p, "normal").to_vec3(), "normal"), unwrap(get(p, "offset").to_scalar(), "offset"));
        default:
            break;
    }
    auto lhs = get(p, "A").to_csg_node();  // operations/mod.rs:47  (`?` -> None)
    if (!lhs) return std::nullopt;
    switch (kind) {  // N4 space transforms: one connection-only input, then constants
        case Kind::Translation: return CSGNode::Translation(*lhs, unwrap(get(p, "offset").to_vec3(), "offset"));
        case Kind::Rotation:
            return CSGNode::Rotation(*lhs, quat_from_axis_angle(unwrap(get(p, "axis").to_vec3(), "axis"), unwrap(get(p, "angle").to_scalar(), "angle")));
        case Kind::Scale: return CSGNode::Scale(*lhs, unwrap(get(p, "factor").to_scalar(), "factor"));
        default: break;
    }
    auto rhs = get(p, "B").to_csg_node();  // operations/mod.rs:48
    if (!rhs) return std::nullopt;
    switch (kind) {
        case Kind::Union: return CSGNode::Union(*lhs, *rhs);
        case Kind::Subtraction: return CSGNode::Subtraction(*lhs, *rhs);
        case Kind::Intersection: return CSGNode::Intersection(*lhs, *rhs);
        default: return CSGNode::SmoothUnion(*lhs, *rhs, unwrap(get(p, "k").to_scalar(), "k"));
    }
}

std::vector<CSGNodeTemplate> CSGNodeTemplate::all() {
    return {{Kind::Sphere}, {Kind::Box}, {Kind::Union}, {Kind::Subtraction}, {Kind::Intersection}, {Kind::SmoothUnion},
            {Kind::Plane}, {Kind::Translation}, {Kind::Rotation}, {Kind::Scale}};
}

}  // namespace csg

ValueType ValueType::Scalar(float x) {
    ValueType v;
    v.tag = Tag::Scalar;
    v.scalar = x;
    return v;
}
ValueType ValueType::Vec3v(const Vec3 &x) {
    ValueType v;
    v.tag = Tag::Vec3;
    v.vec3 = x;
    return v;
}
ValueType ValueType::CSGNode(std::optional<csg::CSGNode> n) {
    ValueType v;
    v.tag = Tag::CSGNode;
    v.node = std::move(n);
    return v;
}
std::optional<float> ValueType::to_scalar() const { return tag == Tag::Scalar ? std::optional<float>(scalar) : std::nullopt; }
std::optional<Vec3> ValueType::to_vec3() const { return tag == Tag::Vec3 ? std::optional<Vec3>(vec3) : std::nullopt; }
std::optional<csg::CSGNode> ValueType::to_csg_node() const { return tag == Tag::CSGNode ? node : std::nullopt; }

}  // namespace ray_marching
