// pts_math.h — minimal host vector/matrix maths for libptscene.
//
// The reference does its host maths with glm 0.9.9.8 (vendored under /root/reference/include/glm).
// libptscene must produce bit-identical vertex data, so each helper below performs the same IEEE
// binary32 operations in the same order as the glm function the reference calls (cited per helper);
// tests/test_scene_vs_ref.py checks the results bit-for-bit against oracle/_ref/libref_host.so.
// Build with -ffp-contract=off.
#pragma once
#include <cmath>

namespace pts {

struct V3 {
    float x, y, z;
};
struct V4 {
    float x, y, z, w;
};
struct M4 {        // column-major, c[col]
    V4 c[4];
};

inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 operator*(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline V4 operator*(V4 a, float s) { return {a.x * s, a.y * s, a.z * s, a.w * s}; }
inline V4 operator*(V4 a, V4 b) { return {a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w}; }
inline V4 operator+(V4 a, V4 b) { return {a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w}; }
inline V4 operator-(V4 a, V4 b) { return {a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w}; }

// glm::dot(vec3): tmp = a*b; tmp.x + tmp.y + tmp.z   (detail/func_geometric.inl compute_dot<vec<3,...>>)
inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
// glm::cross (detail/func_geometric.inl compute_cross)
inline V3 cross(V3 x, V3 y) {
    return {x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y};
}
// glm::normalize = v * inversesqrt(dot(v,v)), inversesqrt(x) = 1/sqrt(x)
inline V3 normalize(V3 v) {
    float inv = 1.0f / std::sqrt(dot(v, v));
    return v * inv;
}
inline float radians(float deg) { return deg * 0.01745329251994329576923690768489f; }

inline M4 identity() {
    return {{{1, 0, 0, 0}, {0, 1, 0, 0}, {0, 0, 1, 0}, {0, 0, 0, 1}}};
}
// mat4 * vec4, glm detail/type_mat4x4.inl:561-572: (m0*v0 + m1*v1) + (m2*v2 + m3*v3)
inline V4 mul(const M4& m, V4 v) {
    V4 add0 = m.c[0] * v.x + m.c[1] * v.y;
    V4 add1 = m.c[2] * v.z + m.c[3] * v.w;
    return add0 + add1;
}
// mat4 * mat4, glm detail/type_mat4x4.inl:630-648: ((A0*b0 + A1*b1) + A2*b2) + A3*b3 per column
inline M4 mul(const M4& a, const M4& b) {
    M4 r;
    for (int j = 0; j < 4; j++)
        r.c[j] = a.c[0] * b.c[j].x + a.c[1] * b.c[j].y + a.c[2] * b.c[j].z + a.c[3] * b.c[j].w;
    return r;
}
// glm::translate, ext/matrix_transform.inl
inline M4 translate(const M4& m, V3 v) {
    M4 r = m;
    r.c[3] = m.c[0] * v.x + m.c[1] * v.y + m.c[2] * v.z + m.c[3];
    return r;
}
// glm::scale
inline M4 scale(const M4& m, V3 v) {
    M4 r;
    r.c[0] = m.c[0] * v.x; r.c[1] = m.c[1] * v.y; r.c[2] = m.c[2] * v.z; r.c[3] = m.c[3];
    return r;
}
// glm::rotate(m, angle, axis)
inline M4 rotate(const M4& m, float angle, V3 v) {
    const float c = std::cos(angle), s = std::sin(angle);
    V3 axis = normalize(v);
    V3 temp = axis * (1.0f - c);                    // (T(1) - c) * axis : scalar*vec, commutative per component
    float R00 = c + temp.x * axis.x, R01 = temp.x * axis.y + s * axis.z, R02 = temp.x * axis.z - s * axis.y;
    float R10 = temp.y * axis.x - s * axis.z, R11 = c + temp.y * axis.y, R12 = temp.y * axis.z + s * axis.x;
    float R20 = temp.z * axis.x + s * axis.y, R21 = temp.z * axis.y - s * axis.x, R22 = c + temp.z * axis.z;
    M4 r;
    r.c[0] = m.c[0] * R00 + m.c[1] * R01 + m.c[2] * R02;
    r.c[1] = m.c[0] * R10 + m.c[1] * R11 + m.c[2] * R12;
    r.c[2] = m.c[0] * R20 + m.c[1] * R21 + m.c[2] * R22;
    r.c[3] = m.c[3];
    return r;
}
// glm::lookAtRH
inline M4 lookAt(V3 eye, V3 center, V3 up) {
    V3 f = normalize(center - eye);
    V3 s = normalize(cross(f, up));
    V3 u = cross(s, f);
    M4 r = identity();
    r.c[0].x = s.x; r.c[1].x = s.y; r.c[2].x = s.z;
    r.c[0].y = u.x; r.c[1].y = u.y; r.c[2].y = u.z;
    r.c[0].z = -f.x; r.c[1].z = -f.y; r.c[2].z = -f.z;
    r.c[3].x = -dot(s, eye); r.c[3].y = -dot(u, eye); r.c[3].z = dot(f, eye);
    return r;
}
// glm::inverse(mat4), detail/func_matrix.inl compute_inverse<4,4>
inline M4 inverse(const M4& mm) {
    auto m = [&](int c, int r) -> float { return (&mm.c[c].x)[r]; };
    float Coef00 = m(2,2) * m(3,3) - m(3,2) * m(2,3);
    float Coef02 = m(1,2) * m(3,3) - m(3,2) * m(1,3);
    float Coef03 = m(1,2) * m(2,3) - m(2,2) * m(1,3);
    float Coef04 = m(2,1) * m(3,3) - m(3,1) * m(2,3);
    float Coef06 = m(1,1) * m(3,3) - m(3,1) * m(1,3);
    float Coef07 = m(1,1) * m(2,3) - m(2,1) * m(1,3);
    float Coef08 = m(2,1) * m(3,2) - m(3,1) * m(2,2);
    float Coef10 = m(1,1) * m(3,2) - m(3,1) * m(1,2);
    float Coef11 = m(1,1) * m(2,2) - m(2,1) * m(1,2);
    float Coef12 = m(2,0) * m(3,3) - m(3,0) * m(2,3);
    float Coef14 = m(1,0) * m(3,3) - m(3,0) * m(1,3);
    float Coef15 = m(1,0) * m(2,3) - m(2,0) * m(1,3);
    float Coef16 = m(2,0) * m(3,2) - m(3,0) * m(2,2);
    float Coef18 = m(1,0) * m(3,2) - m(3,0) * m(1,2);
    float Coef19 = m(1,0) * m(2,2) - m(2,0) * m(1,2);
    float Coef20 = m(2,0) * m(3,1) - m(3,0) * m(2,1);
    float Coef22 = m(1,0) * m(3,1) - m(3,0) * m(1,1);
    float Coef23 = m(1,0) * m(2,1) - m(2,0) * m(1,1);
    V4 Fac0{Coef00, Coef00, Coef02, Coef03}, Fac1{Coef04, Coef04, Coef06, Coef07}, Fac2{Coef08, Coef08, Coef10, Coef11};
    V4 Fac3{Coef12, Coef12, Coef14, Coef15}, Fac4{Coef16, Coef16, Coef18, Coef19}, Fac5{Coef20, Coef20, Coef22, Coef23};
    V4 Vec0{m(1,0), m(0,0), m(0,0), m(0,0)}, Vec1{m(1,1), m(0,1), m(0,1), m(0,1)};
    V4 Vec2{m(1,2), m(0,2), m(0,2), m(0,2)}, Vec3{m(1,3), m(0,3), m(0,3), m(0,3)};
    V4 Inv0 = Vec1 * Fac0 - Vec2 * Fac1 + Vec3 * Fac2;
    V4 Inv1 = Vec0 * Fac0 - Vec2 * Fac3 + Vec3 * Fac4;
    V4 Inv2 = Vec0 * Fac1 - Vec1 * Fac3 + Vec3 * Fac5;
    V4 Inv3 = Vec0 * Fac2 - Vec1 * Fac4 + Vec2 * Fac5;
    V4 SignA{+1, -1, +1, -1}, SignB{-1, +1, -1, +1};
    M4 Inverse{{Inv0 * SignA, Inv1 * SignB, Inv2 * SignA, Inv3 * SignB}};
    V4 Row0{Inverse.c[0].x, Inverse.c[1].x, Inverse.c[2].x, Inverse.c[3].x};
    V4 Dot0 = mm.c[0] * Row0;
    float Dot1 = (Dot0.x + Dot0.y) + (Dot0.z + Dot0.w);
    float OneOverDeterminant = 1.0f / Dot1;
    M4 r;
    for (int j = 0; j < 4; j++) r.c[j] = Inverse.c[j] * OneOverDeterminant;
    return r;
}

}  // namespace pts
