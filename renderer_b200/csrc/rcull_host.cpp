// rcull_host.cpp -- host-side per-view math: the values the
// reference computes once per view per frame in project_camera
// (src/ecs.rs:66-91) and shadow_mapping_mvp_calculation
// (src/renderer/systems/shadow_mapping.rs:314-342).  Planes are INPUTS to the
// frame kernel; this file only mirrors how the reference derives them.
// Column-major 4x4 throughout (nalgebra / GLSL mat4 layout).
#include <cmath>
#include <cstring>

#include "rcull.h"

namespace {

struct M4 {
    float v[16];
    float& at(int r, int c) { return v[c * 4 + r]; }
    float at(int r, int c) const { return v[c * 4 + r]; }
};

M4 zero() {
    M4 m;
    std::memset(m.v, 0, sizeof m.v);
    return m;
}
M4 identity() {
    M4 m = zero();
    m.at(0, 0) = m.at(1, 1) = m.at(2, 2) = m.at(3, 3) = 1.f;
    return m;
}
// column-by-column axpy accumulation, the order nalgebra's small gemm uses
M4 mul(const M4& a, const M4& b) {
    M4 o;
    for (int j = 0; j < 4; ++j)
        for (int i = 0; i < 4; ++i) {
            float acc = a.at(i, 0) * b.at(0, j);
            for (int k = 1; k < 4; ++k) acc = a.at(i, k) * b.at(k, j) + acc;
            o.at(i, j) = acc;
        }
    return o;
}
// glm::perspective_lh_zo(aspect, fovy, near, far)
M4 perspective_lh_zo(float aspect, float fovy, float z_near, float z_far) {
    M4 m = zero();
    const float t = std::tan(fovy / 2.f);
    m.at(0, 0) = 1.f / (aspect * t);
    m.at(1, 1) = 1.f / t;
    m.at(2, 2) = z_far / (z_far - z_near);
    m.at(2, 3) = -(z_far * z_near) / (z_far - z_near);
    m.at(3, 2) = 1.f;
    return m;
}
// UnitQuaternion (i,j,k,w) -> homogeneous rotation
M4 rotation(const float q[4]) {
    const float i = q[0], j = q[1], k = q[2], w = q[3];
    const float ww = w * w, ii = i * i, jj = j * j, kk = k * k;
    const float ij = i * j * 2.f, wk = w * k * 2.f, wj = w * j * 2.f;
    const float ik = i * k * 2.f, jk = j * k * 2.f, wi = w * i * 2.f;
    M4 m = identity();
    m.at(0, 0) = ww + ii - jj - kk; m.at(0, 1) = ij - wk;           m.at(0, 2) = wj + ik;
    m.at(1, 0) = wk + ij;           m.at(1, 1) = ww - ii + jj - kk; m.at(1, 2) = jk - wi;
    m.at(2, 0) = ik - wj;           m.at(2, 1) = wi + jk;           m.at(2, 2) = ww - ii - jj + kk;
    return m;
}
float dot(const float a[3], const float b[3]) { return (a[0] * b[0] + a[1] * b[1]) + a[2] * b[2]; }
void normalize(float v[3]) {
    const float n = std::sqrt(dot(v, v));
    v[0] /= n; v[1] /= n; v[2] /= n;
}
void cross(const float a[3], const float b[3], float o[3]) {
    o[0] = a[1] * b[2] - a[2] * b[1];
    o[1] = a[2] * b[0] - a[0] * b[2];
    o[2] = a[0] * b[1] - a[1] * b[0];
}
// UnitQuaternion * Vector3 as nalgebra 0.29 evaluates it (transform_vector, `rot * v`):
// t = (q.ijk x v) * 2; ((t * w) + (q.ijk x t)) + v
void rotate(const float q[4], const float v[3], float o[3]) {
    float t[3], c[3];
    cross(q, v, t);
    for (int a = 0; a < 3; ++a) t[a] = t[a] * 2.f;
    cross(q, t, c);
    for (int a = 0; a < 3; ++a) o[a] = (t[a] * q[3] + c[a]) + v[a];
}
constexpr float kFov70 = 70.0f * 0.017453292519943295f;  // glm::radians(70)

}  // namespace

extern "C" {

// camera.frustum_planes (src/ecs.rs:82-90): -(row3 + row_i), -(row3 - row_i).
void rcull_frustum_planes(const float pv[16], rcull_view* out) {
    for (int i = 0; i < 3; ++i)
        for (int c = 0; c < 4; ++c) {
            const float r3 = pv[c * 4 + 3], ri = pv[c * 4 + i];
            out->planes[2 * i][c] = -(r3 + ri);
            out->planes[2 * i + 1][c] = -(r3 - ri);
        }
}

// project_camera (src/ecs.rs:66-91): near 0.1, far 100, fov_y 70 degrees;
// view = look_at_lh(position, position + rot * z, rot * y).
void rcull_project_camera(float aspect, const float pos[3], const float rot[4], float proj[16], float view[16],
                          float pv[16], rcull_view* out) {
    const M4 p = perspective_lh_zo(aspect, kFov70, 0.1f, 100.0f);
    const float z[3] = {0, 0, 1}, y[3] = {0, 1, 0};
    float dir[3], f[3], up[3], s[3], u[3];
    rotate(rot, z, dir);
    rotate(rot, y, up);
    for (int a = 0; a < 3; ++a) f[a] = (pos[a] + dir[a]) - pos[a];  // look_at_lh(eye, eye + dir, up): target - eye
    normalize(f);
    cross(up, f, s);
    normalize(s);
    cross(f, s, u);
    M4 v = identity();
    for (int c = 0; c < 3; ++c) {
        v.at(0, c) = s[c];
        v.at(1, c) = u[c];
        v.at(2, c) = f[c];
    }
    v.at(0, 3) = -dot(s, pos);
    v.at(1, 3) = -dot(u, pos);
    v.at(2, 3) = -dot(f, pos);
    const M4 m = mul(p, v);
    if (proj) std::memcpy(proj, p.v, sizeof p.v);
    if (view) std::memcpy(view, v.v, sizeof v.v);
    if (pv) std::memcpy(pv, m.v, sizeof m.v);
    if (out) rcull_frustum_planes(m.v, out);
}

// shadow_mapping_mvp_calculation (shadow_mapping.rs:321-329):
// projection = perspective_lh_zo(1, 70 deg, 10, 400),
// view = translation(rot * (-pos)) * rot.to_homogeneous().
void rcull_light_view(const float pos[3], const float rot[4], float proj[16], float view[16], float pv[16],
                      rcull_view* out) {
    const M4 p = perspective_lh_zo(1.0f, kFov70, 10.0f, 400.0f);
    const M4 r = rotation(rot);
    const float neg[3] = {-pos[0], -pos[1], -pos[2]};
    float t[3];
    rotate(rot, neg, t);
    M4 tm = identity();
    tm.at(0, 3) = t[0];
    tm.at(1, 3) = t[1];
    tm.at(2, 3) = t[2];
    const M4 v = mul(tm, r);
    const M4 m = mul(p, v);
    if (proj) std::memcpy(proj, p.v, sizeof p.v);
    if (view) std::memcpy(view, v.v, sizeof v.v);
    if (pv) std::memcpy(pv, m.v, sizeof m.v);
    if (out) rcull_frustum_planes(m.v, out);
}

}  // extern "C"
