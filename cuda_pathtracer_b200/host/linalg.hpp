// linalg.hpp -- the handful of float vector / 4x4 matrix routines the scene loader needs.
//
// The render is bit-reproducible only if the Geom matrices and the camera handed to the device
// are bit-identical to what the reference's loader computes with GLM 0.9.6 on the host
// (path_tracer/src/scene.cpp:84-96, utilities.cpp:84-92, main.cpp:59-73,121-134).  So every
// routine here performs the SAME float operations in the SAME order as the GLM routine named in
// its comment (plain C++ float arithmetic: compiled for baseline x86-64, where there is no FMA
// instruction to contract into; the Makefile also passes -ffp-contract=off).  Matrices are
// column-major float[16]: element (column c, row r) is m[c*4 + r], as glm::mat4 stores it.
// tests/test_host_cli.py checks the result byte for byte against dumps made by the reference's
// own loader (tests/golden/*.scene.bin).
#pragma once
#include <cmath>

namespace b200host {

struct Vec3 {
    float x, y, z;
};
struct Vec4 {
    float x, y, z, w;
};

inline Vec3 operator+(const Vec3& a, const Vec3& b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline Vec3 operator-(const Vec3& a, const Vec3& b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline Vec3 operator-(const Vec3& a) { return {-a.x, -a.y, -a.z}; }
inline Vec3 operator*(const Vec3& a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline Vec3 operator*(float s, const Vec3& a) { return {s * a.x, s * a.y, s * a.z}; }
inline Vec4 operator+(const Vec4& a, const Vec4& b) { return {a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w}; }
inline Vec4 operator-(const Vec4& a, const Vec4& b) { return {a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w}; }
inline Vec4 operator*(const Vec4& a, const Vec4& b) { return {a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w}; }
inline Vec4 operator*(const Vec4& a, float s) { return {a.x * s, a.y * s, a.z * s, a.w * s}; }

// glm::dot(vec3): tmp = a * b; tmp.x + tmp.y + tmp.z
inline float dot(const Vec3& a, const Vec3& b)
{
    const float tx = a.x * b.x, ty = a.y * b.y, tz = a.z * b.z;
    return tx + ty + tz;
}
// glm::cross
inline Vec3 cross(const Vec3& a, const Vec3& b)
{
    return {a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y};
}
// glm::length = sqrt(dot(v, v)); glm::normalize = v * inversesqrt(dot(v, v)), inversesqrt = 1 / sqrt
inline float length(const Vec3& v) { return std::sqrt(dot(v, v)); }
inline Vec3 normalize(const Vec3& v) { return v * (1.0f / std::sqrt(dot(v, v))); }

struct Mat4 {
    float m[16];
    Vec4 col(int c) const { return {m[c * 4 + 0], m[c * 4 + 1], m[c * 4 + 2], m[c * 4 + 3]}; }
    void set_col(int c, const Vec4& v) { m[c * 4 + 0] = v.x; m[c * 4 + 1] = v.y; m[c * 4 + 2] = v.z; m[c * 4 + 3] = v.w; }
    float at(int c, int r) const { return m[c * 4 + r]; }
};

inline Mat4 identity()
{
    Mat4 r{};
    r.m[0] = r.m[5] = r.m[10] = r.m[15] = 1.0f;
    return r;
}

// glm::translate(m, v): Result[3] = m[0]*v.x + m[1]*v.y + m[2]*v.z + m[3]
inline Mat4 translate(const Mat4& m, const Vec3& v)
{
    Mat4 r = m;
    r.set_col(3, m.col(0) * v.x + m.col(1) * v.y + m.col(2) * v.z + m.col(3));
    return r;
}

// glm::rotate(m, angle, axis) (gtc/matrix_transform.inl)
inline Mat4 rotate(const Mat4& m, float angle, const Vec3& v)
{
    const float c = std::cos(angle), s = std::sin(angle);
    const Vec3 axis = normalize(v);
    const Vec3 temp = (1.0f - c) * axis;
    float R[3][3];
    R[0][0] = c + temp.x * axis.x;
    R[0][1] = 0 + temp.x * axis.y + s * axis.z;
    R[0][2] = 0 + temp.x * axis.z - s * axis.y;
    R[1][0] = 0 + temp.y * axis.x - s * axis.z;
    R[1][1] = c + temp.y * axis.y;
    R[1][2] = 0 + temp.y * axis.z + s * axis.x;
    R[2][0] = 0 + temp.z * axis.x + s * axis.y;
    R[2][1] = 0 + temp.z * axis.y - s * axis.x;
    R[2][2] = c + temp.z * axis.z;
    Mat4 r;
    r.set_col(0, m.col(0) * R[0][0] + m.col(1) * R[0][1] + m.col(2) * R[0][2]);
    r.set_col(1, m.col(0) * R[1][0] + m.col(1) * R[1][1] + m.col(2) * R[1][2]);
    r.set_col(2, m.col(0) * R[2][0] + m.col(1) * R[2][1] + m.col(2) * R[2][2]);
    r.set_col(3, m.col(3));
    return r;
}

// glm::scale(m, v)
inline Mat4 scale(const Mat4& m, const Vec3& v)
{
    Mat4 r;
    r.set_col(0, m.col(0) * v.x);
    r.set_col(1, m.col(1) * v.y);
    r.set_col(2, m.col(2) * v.z);
    r.set_col(3, m.col(3));
    return r;
}

// glm::mat4 * glm::mat4 (detail/type_mat4x4.inl)
inline Mat4 mul(const Mat4& a, const Mat4& b)
{
    Mat4 r;
    for (int c = 0; c < 4; ++c)
        r.set_col(c, a.col(0) * b.at(c, 0) + a.col(1) * b.at(c, 1) + a.col(2) * b.at(c, 2) + a.col(3) * b.at(c, 3));
    return r;
}

// glm::mat4 * glm::vec4: ((m0*v.x + m1*v.y) + (m2*v.z + m3*v.w))
inline Vec4 mul(const Mat4& m, const Vec4& v)
{
    const Vec4 mul0 = m.col(0) * Vec4{v.x, v.x, v.x, v.x};
    const Vec4 mul1 = m.col(1) * Vec4{v.y, v.y, v.y, v.y};
    const Vec4 add0 = mul0 + mul1;
    const Vec4 mul2 = m.col(2) * Vec4{v.z, v.z, v.z, v.z};
    const Vec4 mul3 = m.col(3) * Vec4{v.w, v.w, v.w, v.w};
    const Vec4 add1 = mul2 + mul3;
    return add0 + add1;
}

// glm::inverse(mat4) (detail::compute_inverse, type_mat4x4.inl:37-87)
inline Mat4 inverse(const Mat4& M)
{
    auto m = [&](int c, int r) { return M.at(c, r); };
    const float c00 = m(2, 2) * m(3, 3) - m(3, 2) * m(2, 3);
    const float c02 = m(1, 2) * m(3, 3) - m(3, 2) * m(1, 3);
    const float c03 = m(1, 2) * m(2, 3) - m(2, 2) * m(1, 3);
    const float c04 = m(2, 1) * m(3, 3) - m(3, 1) * m(2, 3);
    const float c06 = m(1, 1) * m(3, 3) - m(3, 1) * m(1, 3);
    const float c07 = m(1, 1) * m(2, 3) - m(2, 1) * m(1, 3);
    const float c08 = m(2, 1) * m(3, 2) - m(3, 1) * m(2, 2);
    const float c10 = m(1, 1) * m(3, 2) - m(3, 1) * m(1, 2);
    const float c11 = m(1, 1) * m(2, 2) - m(2, 1) * m(1, 2);
    const float c12 = m(2, 0) * m(3, 3) - m(3, 0) * m(2, 3);
    const float c14 = m(1, 0) * m(3, 3) - m(3, 0) * m(1, 3);
    const float c15 = m(1, 0) * m(2, 3) - m(2, 0) * m(1, 3);
    const float c16 = m(2, 0) * m(3, 2) - m(3, 0) * m(2, 2);
    const float c18 = m(1, 0) * m(3, 2) - m(3, 0) * m(1, 2);
    const float c19 = m(1, 0) * m(2, 2) - m(2, 0) * m(1, 2);
    const float c20 = m(2, 0) * m(3, 1) - m(3, 0) * m(2, 1);
    const float c22 = m(1, 0) * m(3, 1) - m(3, 0) * m(1, 1);
    const float c23 = m(1, 0) * m(2, 1) - m(2, 0) * m(1, 1);
    const Vec4 f0{c00, c00, c02, c03}, f1{c04, c04, c06, c07}, f2{c08, c08, c10, c11};
    const Vec4 f3{c12, c12, c14, c15}, f4{c16, c16, c18, c19}, f5{c20, c20, c22, c23};
    const Vec4 v0{m(1, 0), m(0, 0), m(0, 0), m(0, 0)}, v1{m(1, 1), m(0, 1), m(0, 1), m(0, 1)};
    const Vec4 v2{m(1, 2), m(0, 2), m(0, 2), m(0, 2)}, v3{m(1, 3), m(0, 3), m(0, 3), m(0, 3)};
    const Vec4 i0 = v1 * f0 - v2 * f1 + v3 * f2;
    const Vec4 i1 = v0 * f0 - v2 * f3 + v3 * f4;
    const Vec4 i2 = v0 * f1 - v1 * f3 + v3 * f5;
    const Vec4 i3 = v0 * f2 - v1 * f4 + v2 * f5;
    const Vec4 sa{+1, -1, +1, -1}, sb{-1, +1, -1, +1};
    Mat4 inv;
    inv.set_col(0, i0 * sa);
    inv.set_col(1, i1 * sb);
    inv.set_col(2, i2 * sa);
    inv.set_col(3, i3 * sb);
    const Vec4 row0{inv.at(0, 0), inv.at(1, 0), inv.at(2, 0), inv.at(3, 0)};
    const Vec4 d0 = M.col(0) * row0;
    const float det = (d0.x + d0.y) + (d0.z + d0.w);
    const float one_over = 1.0f / det;
    Mat4 r;
    for (int c = 0; c < 4; ++c) r.set_col(c, inv.col(c) * one_over);
    return r;
}

// glm::inverseTranspose(mat4) (gtc/matrix_inverse.inl): cofactors divided by the determinant
inline Mat4 inverse_transpose(const Mat4& M)
{
    auto m = [&](int c, int r) { return M.at(c, r); };
    const float s00 = m(2, 2) * m(3, 3) - m(3, 2) * m(2, 3);
    const float s01 = m(2, 1) * m(3, 3) - m(3, 1) * m(2, 3);
    const float s02 = m(2, 1) * m(3, 2) - m(3, 1) * m(2, 2);
    const float s03 = m(2, 0) * m(3, 3) - m(3, 0) * m(2, 3);
    const float s04 = m(2, 0) * m(3, 2) - m(3, 0) * m(2, 2);
    const float s05 = m(2, 0) * m(3, 1) - m(3, 0) * m(2, 1);
    const float s06 = m(1, 2) * m(3, 3) - m(3, 2) * m(1, 3);
    const float s07 = m(1, 1) * m(3, 3) - m(3, 1) * m(1, 3);
    const float s08 = m(1, 1) * m(3, 2) - m(3, 1) * m(1, 2);
    const float s09 = m(1, 0) * m(3, 3) - m(3, 0) * m(1, 3);
    const float s10 = m(1, 0) * m(3, 2) - m(3, 0) * m(1, 2);
    const float s11 = m(1, 1) * m(3, 3) - m(3, 1) * m(1, 3);
    const float s12 = m(1, 0) * m(3, 1) - m(3, 0) * m(1, 1);
    const float s13 = m(1, 2) * m(2, 3) - m(2, 2) * m(1, 3);
    const float s14 = m(1, 1) * m(2, 3) - m(2, 1) * m(1, 3);
    const float s15 = m(1, 1) * m(2, 2) - m(2, 1) * m(1, 2);
    const float s16 = m(1, 0) * m(2, 3) - m(2, 0) * m(1, 3);
    const float s17 = m(1, 0) * m(2, 2) - m(2, 0) * m(1, 2);
    const float s18 = m(1, 0) * m(2, 1) - m(2, 0) * m(1, 1);
    float I[4][4];
    I[0][0] = +(m(1, 1) * s00 - m(1, 2) * s01 + m(1, 3) * s02);
    I[0][1] = -(m(1, 0) * s00 - m(1, 2) * s03 + m(1, 3) * s04);
    I[0][2] = +(m(1, 0) * s01 - m(1, 1) * s03 + m(1, 3) * s05);
    I[0][3] = -(m(1, 0) * s02 - m(1, 1) * s04 + m(1, 2) * s05);
    I[1][0] = -(m(0, 1) * s00 - m(0, 2) * s01 + m(0, 3) * s02);
    I[1][1] = +(m(0, 0) * s00 - m(0, 2) * s03 + m(0, 3) * s04);
    I[1][2] = -(m(0, 0) * s01 - m(0, 1) * s03 + m(0, 3) * s05);
    I[1][3] = +(m(0, 0) * s02 - m(0, 1) * s04 + m(0, 2) * s05);
    I[2][0] = +(m(0, 1) * s06 - m(0, 2) * s07 + m(0, 3) * s08);
    I[2][1] = -(m(0, 0) * s06 - m(0, 2) * s09 + m(0, 3) * s10);
    I[2][2] = +(m(0, 0) * s11 - m(0, 1) * s09 + m(0, 3) * s12);
    I[2][3] = -(m(0, 0) * s08 - m(0, 1) * s10 + m(0, 2) * s12);
    I[3][0] = -(m(0, 1) * s13 - m(0, 2) * s14 + m(0, 3) * s15);
    I[3][1] = +(m(0, 0) * s13 - m(0, 2) * s16 + m(0, 3) * s17);
    I[3][2] = -(m(0, 0) * s14 - m(0, 1) * s16 + m(0, 3) * s18);
    I[3][3] = +(m(0, 0) * s15 - m(0, 1) * s17 + m(0, 2) * s18);
    const float det = +m(0, 0) * I[0][0] + m(0, 1) * I[0][1] + m(0, 2) * I[0][2] + m(0, 3) * I[0][3];
    Mat4 r;
    for (int c = 0; c < 4; ++c)
        for (int k = 0; k < 4; ++k) r.m[c * 4 + k] = I[c][k] / det;
    return r;
}

}  // namespace b200host
