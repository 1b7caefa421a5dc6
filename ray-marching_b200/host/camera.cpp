// camera.cpp — Camera / OrbitCameraController (src/camera.rs) and the projection / inverse math the
// render pass needs (renderer.rs:205-211).  The reference gets this arithmetic from nalgebra 0.32.4,
// which is not under /root/reference; the formulas below restate nalgebra's published algorithms
// in float32 (UnitQuaternion::from_euler_angles, quaternion * vector, Perspective3::new/inverse,
// 4x4 cofactor inverse).  This math is upstream of the C ABI: kernel and oracle both consume the
// resulting 144 bytes, so ULP differences here can change a view, never a parity result.
#include <algorithm>
#include <cmath>

#include "rm_host.hpp"

namespace ray_marching {

Quat quat_from_euler_angles(float roll, float pitch, float yaw) {
    const float sr = std::sin(roll * 0.5f), cr = std::cos(roll * 0.5f);
    const float sp = std::sin(pitch * 0.5f), cp = std::cos(pitch * 0.5f);
    const float sy = std::sin(yaw * 0.5f), cy = std::cos(yaw * 0.5f);
    Quat q;
    q.w = cr * cp * cy + sr * sp * sy;
    q.i = sr * cp * cy - cr * sp * sy;
    q.j = cr * sp * cy + sr * cp * sy;
    q.k = cr * cp * sy - sr * sp * cy;
    return q;
}

static inline Vec3 cross(const Vec3 &a, const Vec3 &b) {
    return {a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]};
}

Vec3 quat_rotate(const Quat &q, const Vec3 &v) {
    // t = 2 * (q.xyz x v);  v' = v + w*t + q.xyz x t
    const Vec3 u{q.i, q.j, q.k};
    Vec3 t = cross(u, v);
    for (float &x : t) x *= 2.0f;
    const Vec3 c = cross(u, t);
    return {v[0] + q.w * t[0] + c[0], v[1] + q.w * t[1] + c[1], v[2] + q.w * t[2] + c[2]};
}

static Mat4 rotation_matrix(const Quat &q) {  // UnitQuaternion::to_rotation_matrix, homogeneous
    const float ww = q.w * q.w, ii = q.i * q.i, jj = q.j * q.j, kk = q.k * q.k;
    const float ij = q.i * q.j * 2.0f, wk = q.w * q.k * 2.0f, wj = q.w * q.j * 2.0f;
    const float ik = q.i * q.k * 2.0f, jk = q.j * q.k * 2.0f, wi = q.w * q.i * 2.0f;
    Mat4 m{};
    // column-major: m[col*4 + row]
    m[0] = ww + ii - jj - kk; m[4] = ij - wk;           m[8] = wj + ik;            m[12] = 0;
    m[1] = wk + ij;           m[5] = ww - ii + jj - kk; m[9] = jk - wi;            m[13] = 0;
    m[2] = ik - wj;           m[6] = wi + jk;           m[10] = ww - ii - jj + kk; m[14] = 0;
    m[3] = 0;                 m[7] = 0;                 m[11] = 0;                 m[15] = 1;
    return m;
}

Quat quat_from_axis_angle(const Vec3 &axis, float angle) {
    const float n = std::sqrt(axis[0] * axis[0] + axis[1] * axis[1] + axis[2] * axis[2]);
    if (!(n > 0.0f)) return Quat{};  // no axis: identity
    const float s = std::sin(angle * 0.5f) / n;
    return Quat{std::cos(angle * 0.5f), axis[0] * s, axis[1] * s, axis[2] * s};
}

std::array<float, 9> quat_to_matrix3(const Quat &q) {
    const Mat4 m = rotation_matrix(q);  // column-major 4x4
    return {m[0], m[4], m[8], m[1], m[5], m[9], m[2], m[6], m[10]};
}

Mat4 Camera::view() const {
    // rotation.inverse() * Translation3::from(-position)  (camera.rs:10-12) as a homogeneous matrix:
    // [ R^T | R^T * (-position) ]
    const Quat inv{rotation.w, -rotation.i, -rotation.j, -rotation.k};
    Mat4 m = rotation_matrix(inv);
    const Vec3 t = quat_rotate(inv, {-position[0], -position[1], -position[2]});
    m[12] = t[0]; m[13] = t[1]; m[14] = t[2];
    return m;
}

OrbitCameraController::OrbitCameraController(const Vec3 &target_, float radius_) : target(target_), radius(radius_) {}

Quat OrbitCameraController::rotation() const { return quat_from_euler_angles(-pitch, -yaw, 0.0f); }  // camera.rs:53

Camera OrbitCameraController::camera() const {
    const Quat rot = rotation();
    const Vec3 z = quat_rotate(rot, {0.0f, 0.0f, 1.0f});  // rotation * Vector3::z()
    Camera c;
    c.position = {target[0] + z[0] * radius, target[1] + z[1] * radius, target[2] + z[2] * radius};  // camera.rs:58
    c.rotation = rot;
    return c;
}

void OrbitCameraController::update(const OrbitCameraControllerEvent &e) {
    switch (e.tag) {
        case OrbitCameraControllerEvent::Tag::Pan: {  // camera.rs:64-70
            const Vec3 right = quat_rotate(rotation(), {1.0f, 0.0f, 0.0f});
            const Vec3 up = quat_rotate(rotation(), {0.0f, 1.0f, 0.0f});
            for (int i = 0; i < 3; ++i) target[i] += (right[i] * -e.x + up[i] * e.y) * pan_speed;
            break;
        }
        case OrbitCameraControllerEvent::Tag::Orbit:  // camera.rs:71-78
            yaw += e.x * yaw_speed;
            pitch += e.y * pitch_speed;
            pitch = std::min(std::max(pitch, -1.5f), 1.5f);
            break;
        case OrbitCameraControllerEvent::Tag::Dolly:  // camera.rs:79-82
            radius += e.x * dolly_speed * radius;
            radius = std::max(radius, 0.1f);
            break;
    }
}

Mat4 perspective_inverse(float aspect, float fovy, float znear, float zfar) {
    // Perspective3::new (OpenGL-style -1..1 depth) then Perspective3::inverse (closed form)
    const float t = std::tan(fovy / 2.0f);
    const float m11 = 1.0f / (aspect * t), m22 = 1.0f / t;
    const float m33 = (zfar + znear) / (znear - zfar), m34 = zfar * znear * 2.0f / (znear - zfar);
    Mat4 r{};
    r[0] = 1.0f / m11;       // (0,0)
    r[5] = 1.0f / m22;       // (1,1)
    r[14] = -1.0f;           // (2,3)
    r[11] = 1.0f / m34;      // (3,2)
    r[15] = m33 / m34;       // (3,3)
    return r;
}

Mat4 mat4_inverse(const Mat4 &m) {  // general 4x4 inverse by cofactors (what Affine3::inverse falls back to)
    Mat4 inv;
    inv[0] = m[5] * m[10] * m[15] - m[5] * m[11] * m[14] - m[9] * m[6] * m[15] + m[9] * m[7] * m[14] + m[13] * m[6] * m[11] - m[13] * m[7] * m[10];
    inv[4] = -m[4] * m[10] * m[15] + m[4] * m[11] * m[14] + m[8] * m[6] * m[15] - m[8] * m[7] * m[14] - m[12] * m[6] * m[11] + m[12] * m[7] * m[10];
    inv[8] = m[4] * m[9] * m[15] - m[4] * m[11] * m[13] - m[8] * m[5] * m[15] + m[8] * m[7] * m[13] + m[12] * m[5] * m[11] - m[12] * m[7] * m[9];
    inv[12] = -m[4] * m[9] * m[14] + m[4] * m[10] * m[13] + m[8] * m[5] * m[14] - m[8] * m[6] * m[13] - m[12] * m[5] * m[10] + m[12] * m[6] * m[9];
    inv[1] = -m[1] * m[10] * m[15] + m[1] * m[11] * m[14] + m[9] * m[2] * m[15] - m[9] * m[3] * m[14] - m[13] * m[2] * m[11] + m[13] * m[3] * m[10];
    inv[5] = m[0] * m[10] * m[15] - m[0] * m[11] * m[14] - m[8] * m[2] * m[15] + m[8] * m[3] * m[14] + m[12] * m[2] * m[11] - m[12] * m[3] * m[10];
    inv[9] = -m[0] * m[9] * m[15] + m[0] * m[11] * m[13] + m[8] * m[1] * m[15] - m[8] * m[3] * m[13] - m[12] * m[1] * m[11] + m[12] * m[3] * m[9];
    inv[13] = m[0] * m[9] * m[14] - m[0] * m[10] * m[13] - m[8] * m[1] * m[14] + m[8] * m[2] * m[13] + m[12] * m[1] * m[10] - m[12] * m[2] * m[9];
    inv[2] = m[1] * m[6] * m[15] - m[1] * m[7] * m[14] - m[5] * m[2] * m[15] + m[5] * m[3] * m[14] + m[13] * m[2] * m[7] - m[13] * m[3] * m[6];
    inv[6] = -m[0] * m[6] * m[15] + m[0] * m[7] * m[14] + m[4] * m[2] * m[15] - m[4] * m[3] * m[14] - m[12] * m[2] * m[7] + m[12] * m[3] * m[6];
    inv[10] = m[0] * m[5] * m[15] - m[0] * m[7] * m[13] - m[4] * m[1] * m[15] + m[4] * m[3] * m[13] + m[12] * m[1] * m[7] - m[12] * m[3] * m[5];
    inv[14] = -m[0] * m[5] * m[14] + m[0] * m[6] * m[13] + m[4] * m[1] * m[14] - m[4] * m[2] * m[13] - m[12] * m[1] * m[6] + m[12] * m[2] * m[5];
    inv[3] = -m[1] * m[6] * m[11] + m[1] * m[7] * m[10] + m[5] * m[2] * m[11] - m[5] * m[3] * m[10] - m[9] * m[2] * m[7] + m[9] * m[3] * m[6];
    inv[7] = m[0] * m[6] * m[11] - m[0] * m[7] * m[10] - m[4] * m[2] * m[11] + m[4] * m[3] * m[10] + m[8] * m[2] * m[7] - m[8] * m[3] * m[6];
    inv[11] = -m[0] * m[5] * m[11] + m[0] * m[7] * m[9] + m[4] * m[1] * m[11] - m[4] * m[3] * m[9] - m[8] * m[1] * m[7] + m[8] * m[3] * m[5];
    inv[15] = m[0] * m[5] * m[10] - m[0] * m[6] * m[9] - m[4] * m[1] * m[10] + m[4] * m[2] * m[9] + m[8] * m[1] * m[6] - m[8] * m[2] * m[5];
    const float det = m[0] * inv[0] + m[1] * inv[4] + m[2] * inv[8] + m[3] * inv[12];
    const float inv_det = 1.0f / det;
    for (float &x : inv) x *= inv_det;
    return inv;
}

}  // namespace ray_marching
