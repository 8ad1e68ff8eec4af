// Math.hpp — the handful of glm 0.9.9.9 functions the reference's Camera / Raytracer::update use, restated with
// the same FP32 evaluation order so the 256-byte uniform block comes out bit-identical to the reference's
// (checked against the reference's own Camera.cpp compiled from /root/reference — tests/golden/camera_*.json).
//   glm::perspective (RH, NO)   submodules/glm/glm/ext/matrix_clip_space.inl:249-262
//   glm::lookAt (RH)            submodules/glm/glm/ext/matrix_transform.inl:99-119
//   glm::yawPitchRoll           submodules/glm/glm/gtx/euler_angles.inl:605-636
//   glm::inverse(mat4)          submodules/glm/glm/detail/func_matrix.inl:347-404
//   mat4 * vec4                 submodules/glm/glm/detail/type_mat4x4.inl:561-573
// vec3 / vec4 / mat4 are layout-compatible with glm::vec3 / glm::vec4 / glm::mat4 (column-major).
#pragma once

#include <cmath>

namespace vkrt
{
struct vec3
{
    float x = 0.0f, y = 0.0f, z = 0.0f;
    vec3() = default;
    vec3(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
};
struct vec4
{
    float x = 0.0f, y = 0.0f, z = 0.0f, w = 0.0f;
    vec4() = default;
    vec4(float x_, float y_, float z_, float w_) : x(x_), y(y_), z(z_), w(w_) {}
    float& operator[](int i) { return (&x)[i]; }
    const float& operator[](int i) const { return (&x)[i]; }
};
struct mat4
{
    vec4 col[4];
    mat4() = default;
    explicit mat4(float d)
    {
        col[0] = vec4(d, 0, 0, 0), col[1] = vec4(0, d, 0, 0), col[2] = vec4(0, 0, d, 0), col[3] = vec4(0, 0, 0, d);
    }
    vec4& operator[](int i) { return col[i]; }
    const vec4& operator[](int i) const { return col[i]; }
};

inline vec3 operator+(vec3 a, vec3 b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline vec3 operator-(vec3 a, vec3 b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline vec3 operator-(vec3 a) { return vec3(-a.x, -a.y, -a.z); }
inline vec3 operator*(vec3 a, float s) { return vec3(a.x * s, a.y * s, a.z * s); }
inline vec3& operator+=(vec3& a, vec3 b)
{
    a = a + b;
    return a;
}
inline vec4 operator*(vec4 a, vec4 b) { return vec4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }
inline vec4 operator*(vec4 a, float s) { return vec4(a.x * s, a.y * s, a.z * s, a.w * s); }
inline vec4 operator+(vec4 a, vec4 b) { return vec4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
inline vec4 operator-(vec4 a, vec4 b) { return vec4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }

inline float dot(vec3 a, vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline vec3 cross(vec3 x, vec3 y) { return vec3(x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y); }
inline vec3 normalize(vec3 v) { return v * (1.0f / std::sqrt(dot(v, v))); }

inline vec4 operator*(const mat4& m, vec4 v)
{
    const vec4 mul0 = m[0] * v.x, mul1 = m[1] * v.y;
    const vec4 add0 = mul0 + mul1;
    const vec4 mul2 = m[2] * v.z, mul3 = m[3] * v.w;
    const vec4 add1 = mul2 + mul3;
    return add0 + add1;
}

// glm mat4 * mat4: every result column is m1 * (column of m2), terms added left to right
inline mat4 operator*(const mat4& m1, const mat4& m2)
{
    mat4 r;
    for (int j = 0; j < 4; ++j)
    {
        r[j] = ((m1[0] * m2[j].x + m1[1] * m2[j].y) + m1[2] * m2[j].z) + m1[3] * m2[j].w;
    }
    return r;
}

// glm::scale(vec3) from <glm/gtx/transform.hpp>: identity with the diagonal scaled
inline mat4 scale(vec3 v)
{
    mat4 r(1.0f);
    r[0] = r[0] * v.x, r[1] = r[1] * v.y, r[2] = r[2] * v.z;
    return r;
}

inline mat4 perspective(float fovy, float aspect, float zNear, float zFar)
{
    const float tanHalfFovy = std::tan(fovy / 2.0f);
    mat4 r(0.0f);
    r[0][0] = 1.0f / (aspect * tanHalfFovy);
    r[1][1] = 1.0f / (tanHalfFovy);
    r[2][2] = -(zFar + zNear) / (zFar - zNear);
    r[2][3] = -1.0f;
    r[3][2] = -(2.0f * zFar * zNear) / (zFar - zNear);
    return r;
}

inline mat4 lookAt(vec3 eye, vec3 center, vec3 up)
{
    const vec3 f(normalize(center - eye));
    const vec3 s(normalize(cross(f, up)));
    const vec3 u(cross(s, f));
    mat4 r(1.0f);
    r[0][0] = s.x, r[1][0] = s.y, r[2][0] = s.z;
    r[0][1] = u.x, r[1][1] = u.y, r[2][1] = u.z;
    r[0][2] = -f.x, r[1][2] = -f.y, r[2][2] = -f.z;
    r[3][0] = -dot(s, eye), r[3][1] = -dot(u, eye), r[3][2] = dot(f, eye);
    return r;
}

inline mat4 yawPitchRoll(float yaw, float pitch, float roll)
{
    const float ch = std::cos(yaw), sh = std::sin(yaw);
    const float cp = std::cos(pitch), sp = std::sin(pitch);
    const float cb = std::cos(roll), sb = std::sin(roll);
    mat4 r;
    r[0][0] = ch * cb + sh * sp * sb;
    r[0][1] = sb * cp;
    r[0][2] = -sh * cb + ch * sp * sb;
    r[0][3] = 0.0f;
    r[1][0] = -ch * sb + sh * sp * cb;
    r[1][1] = cb * cp;
    r[1][2] = sb * sh + ch * sp * cb;
    r[1][3] = 0.0f;
    r[2][0] = sh * cp;
    r[2][1] = -sp;
    r[2][2] = ch * cp;
    r[2][3] = 0.0f;
    r[3][0] = 0.0f, r[3][1] = 0.0f, r[3][2] = 0.0f, r[3][3] = 1.0f;
    return r;
}

inline mat4 inverse(const mat4& m)
{
    const float c00 = m[2][2] * m[3][3] - m[3][2] * m[2][3];
    const float c02 = m[1][2] * m[3][3] - m[3][2] * m[1][3];
    const float c03 = m[1][2] * m[2][3] - m[2][2] * m[1][3];
    const float c04 = m[2][1] * m[3][3] - m[3][1] * m[2][3];
    const float c06 = m[1][1] * m[3][3] - m[3][1] * m[1][3];
    const float c07 = m[1][1] * m[2][3] - m[2][1] * m[1][3];
    const float c08 = m[2][1] * m[3][2] - m[3][1] * m[2][2];
    const float c10 = m[1][1] * m[3][2] - m[3][1] * m[1][2];
    const float c11 = m[1][1] * m[2][2] - m[2][1] * m[1][2];
    const float c12 = m[2][0] * m[3][3] - m[3][0] * m[2][3];
    const float c14 = m[1][0] * m[3][3] - m[3][0] * m[1][3];
    const float c15 = m[1][0] * m[2][3] - m[2][0] * m[1][3];
    const float c16 = m[2][0] * m[3][2] - m[3][0] * m[2][2];
    const float c18 = m[1][0] * m[3][2] - m[3][0] * m[1][2];
    const float c19 = m[1][0] * m[2][2] - m[2][0] * m[1][2];
    const float c20 = m[2][0] * m[3][1] - m[3][0] * m[2][1];
    const float c22 = m[1][0] * m[3][1] - m[3][0] * m[1][1];
    const float c23 = m[1][0] * m[2][1] - m[2][0] * m[1][1];

    const vec4 fac0(c00, c00, c02, c03), fac1(c04, c04, c06, c07), fac2(c08, c08, c10, c11);
    const vec4 fac3(c12, c12, c14, c15), fac4(c16, c16, c18, c19), fac5(c20, c20, c22, c23);
    const vec4 v0(m[1][0], m[0][0], m[0][0], m[0][0]), v1(m[1][1], m[0][1], m[0][1], m[0][1]);
    const vec4 v2(m[1][2], m[0][2], m[0][2], m[0][2]), v3(m[1][3], m[0][3], m[0][3], m[0][3]);

    const vec4 inv0(v1 * fac0 - v2 * fac1 + v3 * fac2);
    const vec4 inv1(v0 * fac0 - v2 * fac3 + v3 * fac4);
    const vec4 inv2(v0 * fac1 - v1 * fac3 + v3 * fac5);
    const vec4 inv3(v0 * fac2 - v1 * fac4 + v2 * fac5);
    const vec4 signA(+1, -1, +1, -1), signB(-1, +1, -1, +1);
    mat4 inv;
    inv[0] = inv0 * signA, inv[1] = inv1 * signB, inv[2] = inv2 * signA, inv[3] = inv3 * signB;
    const vec4 row0(inv[0][0], inv[1][0], inv[2][0], inv[3][0]);
    const vec4 dot0(m[0] * row0);
    const float dot1 = (dot0.x + dot0.y) + (dot0.z + dot0.w);
    const float oneOverDet = 1.0f / dot1;
    mat4 out;
    for (int i = 0; i < 4; ++i)
    {
        out[i] = inv[i] * oneOverDet;
    }
    return out;
}
} // namespace vkrt
