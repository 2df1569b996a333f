// vk_math.h — the few glm 0.9.9.7 operations the engine's host path uses, restated with the same
// association order so the matrices fed to the kernels match what the reference's CPU code computes
// (GLM_FORCE_DEPTH_ZERO_TO_ONE, right-handed: src/CMakeLists.txt:24).  Column-major: m.c[col][row].
// Compile with -ffp-contract=off.
#pragma once
#include <cmath>
#include <cstring>

namespace vkm {

struct vec3 { float x = 0, y = 0, z = 0; };
struct vec4 { float x = 0, y = 0, z = 0, w = 0; };
struct quat { float w = 1, x = 0, y = 0, z = 0; };
struct mat4 {
    float c[4][4];
    mat4() { std::memset(c, 0, sizeof(c)); }
    explicit mat4(float d) { std::memset(c, 0, sizeof(c)); c[0][0] = c[1][1] = c[2][2] = c[3][3] = d; }
    const float* data() const { return &c[0][0]; }
    float* data() { return &c[0][0]; }
};

inline vec3 operator+(vec3 a, vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline vec3 operator-(vec3 a, vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline vec3 operator-(vec3 a) { return {-a.x, -a.y, -a.z}; }
inline vec3 operator*(vec3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline vec3 operator/(vec3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
inline vec4 operator+(vec4 a, vec4 b) { return {a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w}; }
inline vec4 operator-(vec4 a, vec4 b) { return {a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w}; }
inline vec4 operator/(vec4 a, float s) { return {a.x / s, a.y / s, a.z / s, a.w / s}; }
inline vec3 xyz(vec4 v) { return {v.x, v.y, v.z}; }

// glm::dot(vec3): (a.x*b.x + a.y*b.y) + a.z*b.z   (detail/func_geometric.inl compute_dot<vec<3>>)
inline float dot(vec3 a, vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline float length(vec3 a) { return std::sqrt(dot(a, a)); }
// glm::normalize = v * inversesqrt(dot(v,v)), inversesqrt(x) = 1/sqrt(x)
inline vec3 normalize(vec3 a) { return a * (1.0f / std::sqrt(dot(a, a))); }
inline vec3 cross(vec3 a, vec3 b) { return {a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y}; }
inline vec3 min3(vec3 a, vec3 b) { return {std::fmin(a.x, b.x), std::fmin(a.y, b.y), std::fmin(a.z, b.z)}; }
inline vec3 max3(vec3 a, vec3 b) { return {std::fmax(a.x, b.x), std::fmax(a.y, b.y), std::fmax(a.z, b.z)}; }

// mat4 * vec4: (m0*x + m1*y) + (m2*z + m3*w)   (detail/type_mat4x4.inl:561-572)
inline vec4 operator*(const mat4& m, vec4 v) {
    vec4 r;
    float* o = &r.x;
    for (int i = 0; i < 4; ++i) o[i] = (m.c[0][i] * v.x + m.c[1][i] * v.y) + (m.c[2][i] * v.z + m.c[3][i] * v.w);
    return r;
}
// mat4 * mat4: ((A0*b0 + A1*b1) + A2*b2) + A3*b3   (detail/type_mat4x4.inl:606-625)
inline mat4 operator*(const mat4& a, const mat4& b) {
    mat4 r;
    for (int j = 0; j < 4; ++j)
        for (int i = 0; i < 4; ++i)
            r.c[j][i] = ((a.c[0][i] * b.c[j][0] + a.c[1][i] * b.c[j][1]) + a.c[2][i] * b.c[j][2]) + a.c[3][i] * b.c[j][3];
    return r;
}
inline mat4 transpose(const mat4& m) {
    mat4 r;
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) r.c[i][j] = m.c[j][i];
    return r;
}
inline vec4 column(const mat4& m, int j) { return {m.c[j][0], m.c[j][1], m.c[j][2], m.c[j][3]}; }

// glm::translate(mat4(1), v): Result[3] = m[0]*v.x + m[1]*v.y + m[2]*v.z + m[3]  (ext/matrix_transform.inl:6-11)
inline mat4 translate(vec3 v) { mat4 r(1.0f); r.c[3][0] = v.x; r.c[3][1] = v.y; r.c[3][2] = v.z; return r; }
inline mat4 scale(vec3 v) { mat4 r(1.0f); r.c[0][0] = v.x; r.c[1][1] = v.y; r.c[2][2] = v.z; return r; }

// glm::angleAxis (ext/quaternion_trigonometric.inl:27-33)
inline quat angleAxis(float angle, vec3 v) {
    const float s = std::sin(angle * 0.5f);
    return {std::cos(angle * 0.5f), v.x * s, v.y * s, v.z * s};
}
// glm::mat4_cast / toMat4 (gtc/quaternion.inl mat3_cast)
inline mat4 toMat4(quat q) {
    mat4 r(1.0f);
    const float qxx = q.x * q.x, qyy = q.y * q.y, qzz = q.z * q.z, qxz = q.x * q.z, qxy = q.x * q.y, qyz = q.y * q.z;
    const float qwx = q.w * q.x, qwy = q.w * q.y, qwz = q.w * q.z;
    r.c[0][0] = 1.0f - 2.0f * (qyy + qzz); r.c[0][1] = 2.0f * (qxy + qwz); r.c[0][2] = 2.0f * (qxz - qwy);
    r.c[1][0] = 2.0f * (qxy - qwz); r.c[1][1] = 1.0f - 2.0f * (qxx + qzz); r.c[1][2] = 2.0f * (qyz + qwx);
    r.c[2][0] = 2.0f * (qxz + qwy); r.c[2][1] = 2.0f * (qyz - qwx); r.c[2][2] = 1.0f - 2.0f * (qxx + qyy);
    return r;
}

// glm::inverse(mat4) — detail/func_matrix.inl:294-350, same cofactor grouping.
inline mat4 inverse(const mat4& mm) {
    const float (*m)[4] = mm.c;
    const float C00 = m[2][2] * m[3][3] - m[3][2] * m[2][3], C02 = m[1][2] * m[3][3] - m[3][2] * m[1][3], C03 = m[1][2] * m[2][3] - m[2][2] * m[1][3];
    const float C04 = m[2][1] * m[3][3] - m[3][1] * m[2][3], C06 = m[1][1] * m[3][3] - m[3][1] * m[1][3], C07 = m[1][1] * m[2][3] - m[2][1] * m[1][3];
    const float C08 = m[2][1] * m[3][2] - m[3][1] * m[2][2], C10 = m[1][1] * m[3][2] - m[3][1] * m[1][2], C11 = m[1][1] * m[2][2] - m[2][1] * m[1][2];
    const float C12 = m[2][0] * m[3][3] - m[3][0] * m[2][3], C14 = m[1][0] * m[3][3] - m[3][0] * m[1][3], C15 = m[1][0] * m[2][3] - m[2][0] * m[1][3];
    const float C16 = m[2][0] * m[3][2] - m[3][0] * m[2][2], C18 = m[1][0] * m[3][2] - m[3][0] * m[1][2], C19 = m[1][0] * m[2][2] - m[2][0] * m[1][2];
    const float C20 = m[2][0] * m[3][1] - m[3][0] * m[2][1], C22 = m[1][0] * m[3][1] - m[3][0] * m[1][1], C23 = m[1][0] * m[2][1] - m[2][0] * m[1][1];
    const float F0[4] = {C00, C00, C02, C03}, F1[4] = {C04, C04, C06, C07}, F2[4] = {C08, C08, C10, C11};
    const float F3[4] = {C12, C12, C14, C15}, F4[4] = {C16, C16, C18, C19}, F5[4] = {C20, C20, C22, C23};
    const float V0[4] = {m[1][0], m[0][0], m[0][0], m[0][0]}, V1[4] = {m[1][1], m[0][1], m[0][1], m[0][1]};
    const float V2[4] = {m[1][2], m[0][2], m[0][2], m[0][2]}, V3[4] = {m[1][3], m[0][3], m[0][3], m[0][3]};
    const float SA[4] = {1, -1, 1, -1}, SB[4] = {-1, 1, -1, 1};
    mat4 inv;
    for (int i = 0; i < 4; ++i) {
        inv.c[0][i] = (V1[i] * F0[i] - V2[i] * F1[i] + V3[i] * F2[i]) * SA[i];
        inv.c[1][i] = (V0[i] * F0[i] - V2[i] * F3[i] + V3[i] * F4[i]) * SB[i];
        inv.c[2][i] = (V0[i] * F1[i] - V1[i] * F3[i] + V3[i] * F5[i]) * SA[i];
        inv.c[3][i] = (V0[i] * F2[i] - V1[i] * F4[i] + V2[i] * F5[i]) * SB[i];
    }
    const float d0 = m[0][0] * inv.c[0][0], d1 = m[0][1] * inv.c[1][0], d2 = m[0][2] * inv.c[2][0], d3 = m[0][3] * inv.c[3][0];
    const float ood = 1.0f / ((d0 + d1) + (d2 + d3));
    for (int j = 0; j < 4; ++j) for (int i = 0; i < 4; ++i) inv.c[j][i] *= ood;
    return inv;
}

inline float radians(float deg) { return deg * 0.01745329251994329576923690768489f; }

// glm::perspectiveRH_ZO (ext/matrix_clip_space.inl:233-248)
inline mat4 perspectiveRH_ZO(float fovy, float aspect, float zNear, float zFar) {
    const float t = std::tan(fovy / 2.0f);
    mat4 r;
    r.c[0][0] = 1.0f / (aspect * t);
    r.c[1][1] = 1.0f / t;
    r.c[2][2] = zFar / (zNear - zFar);
    r.c[2][3] = -1.0f;
    r.c[3][2] = -(zFar * zNear) / (zFar - zNear);
    return r;
}
// glm::orthoRH_ZO (ext/matrix_clip_space.inl:38-48)
inline mat4 orthoRH_ZO(float l, float r_, float b, float t, float n, float f) {
    mat4 r(1.0f);
    r.c[0][0] = 2.0f / (r_ - l);
    r.c[1][1] = 2.0f / (t - b);
    r.c[2][2] = -1.0f / (f - n);
    r.c[3][0] = -(r_ + l) / (r_ - l);
    r.c[3][1] = -(t + b) / (t - b);
    r.c[3][2] = -n / (f - n);
    return r;
}
// glm::lookAtRH (ext/matrix_transform.inl:98-120)
inline mat4 lookAtRH(vec3 eye, vec3 center, vec3 up) {
    const vec3 f = normalize(center - eye);
    const vec3 s = normalize(cross(f, up));
    const vec3 u = cross(s, f);
    mat4 r(1.0f);
    r.c[0][0] = s.x; r.c[1][0] = s.y; r.c[2][0] = s.z;
    r.c[0][1] = u.x; r.c[1][1] = u.y; r.c[2][1] = u.z;
    r.c[0][2] = -f.x; r.c[1][2] = -f.y; r.c[2][2] = -f.z;
    r.c[3][0] = -dot(s, eye); r.c[3][1] = -dot(u, eye); r.c[3][2] = dot(f, eye);
    return r;
}

}  // namespace vkm
