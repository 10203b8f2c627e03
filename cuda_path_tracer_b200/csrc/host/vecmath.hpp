// vecmath.hpp -- host float vector/matrix helpers whose OPERATION ORDER follows GLM 0.9.6.3
// (vendored by the reference under external/include/glm), because the loader's vertex
// positions, normals, light CDF and camera basis are part of the bit-exact contract
// (triangle order + positions decide the LBVH).
//   dot      : (a.x*b.x + a.y*b.y) + a.z*b.z           detail/func_geometric.inl:65-72
//   cross    : (x.y*y.z - y.y*x.z, ...)                detail/func_geometric.inl:134-142
//   normalize: v * (1 / sqrt(dot(v,v)))                detail/func_geometric.inl:154-159
// Compile this code with -ffp-contract=off (the reference's host code is plain x86-64 g++/MSVC
// code without FMA contraction).
#pragma once
#include <cmath>

namespace ptb {

struct V2 { float x = 0, y = 0; };
struct V3 {
    float x = 0, y = 0, z = 0;
    V3() {}
    V3(float a, float b, float c) : x(a), y(b), z(c) {}
};
struct V4 {
    float x = 0, y = 0, z = 0, w = 0;
    V4() {}
    V4(float a, float b, float c, float d) : x(a), y(b), z(c), w(d) {}
};

inline V3 operator+(V3 a, V3 b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 operator-(V3 a, V3 b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 operator*(V3 a, float s) { return V3(a.x * s, a.y * s, a.z * s); }
inline V3 operator*(float s, V3 a) { return V3(s * a.x, s * a.y, s * a.z); }
inline V4 operator+(V4 a, V4 b) { return V4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
inline V4 operator*(V4 a, float s) { return V4(a.x * s, a.y * s, a.z * s, a.w * s); }
inline V4 operator*(V4 a, V4 b) { return V4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }

inline float dot(V3 a, V3 b) {
    float tx = a.x * b.x, ty = a.y * b.y, tz = a.z * b.z;
    return (tx + ty) + tz;
}
inline V3 cross(V3 x, V3 y) {
    return V3(x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y);
}
inline float length(V3 v) { return std::sqrt(dot(v, v)); }
inline V3 normalize(V3 v) { return v * (1.0f / std::sqrt(dot(v, v))); }

// Column-major 4x4, c[i] = column i (GLM convention, m[col][row]).
struct M4 {
    V4 c[4];
    static M4 identity() {
        M4 m;
        m.c[0] = V4(1, 0, 0, 0); m.c[1] = V4(0, 1, 0, 0); m.c[2] = V4(0, 0, 1, 0); m.c[3] = V4(0, 0, 0, 1);
        return m;
    }
    float at(int col, int row) const {
        const V4& v = c[col];
        return row == 0 ? v.x : row == 1 ? v.y : row == 2 ? v.z : v.w;
    }
};

// gtc/matrix_transform.inl:40-49
inline M4 translate(const M4& m, V3 v) {
    M4 r = m;
    r.c[3] = m.c[0] * v.x + m.c[1] * v.y + m.c[2] * v.z + m.c[3];
    return r;
}
// gtc/matrix_transform.inl:52-85 (axis is already unit length at every call site:
// src/common/utilities.cpp:88-90 passes (1,0,0), (0,1,0), (0,0,1); normalize kept for order)
inline M4 rotate(const M4& m, float angle, V3 v) {
    const float a = angle;
    const float c = std::cos(a);
    const float s = std::sin(a);
    V3 axis = normalize(v);
    V3 temp = (1.0f - c) * axis;
    float R00 = c + temp.x * axis.x;
    float R01 = 0 + temp.x * axis.y + s * axis.z;
    float R02 = 0 + temp.x * axis.z - s * axis.y;
    float R10 = 0 + temp.y * axis.x - s * axis.z;
    float R11 = c + temp.y * axis.y;
    float R12 = 0 + temp.y * axis.z + s * axis.x;
    float R20 = 0 + temp.z * axis.x + s * axis.y;
    float R21 = 0 + temp.z * axis.y - s * axis.x;
    float R22 = c + temp.z * axis.z;
    M4 r;
    r.c[0] = m.c[0] * R00 + m.c[1] * R01 + m.c[2] * R02;
    r.c[1] = m.c[0] * R10 + m.c[1] * R11 + m.c[2] * R12;
    r.c[2] = m.c[0] * R20 + m.c[1] * R21 + m.c[2] * R22;
    r.c[3] = m.c[3];
    return r;
}
// gtc/matrix_transform.inl:122-134
inline M4 scale(const M4& m, V3 v) {
    M4 r;
    r.c[0] = m.c[0] * v.x; r.c[1] = m.c[1] * v.y; r.c[2] = m.c[2] * v.z; r.c[3] = m.c[3];
    return r;
}
// detail/type_mat4x4.inl:686-704
inline M4 operator*(const M4& a, const M4& b) {
    M4 r;
    for (int i = 0; i < 4; i++)
        r.c[i] = a.c[0] * b.c[i].x + a.c[1] * b.c[i].y + a.c[2] * b.c[i].z + a.c[3] * b.c[i].w;
    return r;
}
// detail/type_mat4x4.inl:592-629 : (m0*v0 + m1*v1) + (m2*v2 + m3*v3)
inline V4 operator*(const M4& m, V4 v) {
    V4 add0 = m.c[0] * V4(v.x, v.x, v.x, v.x) + m.c[1] * V4(v.y, v.y, v.y, v.y);
    V4 add1 = m.c[2] * V4(v.z, v.z, v.z, v.z) + m.c[3] * V4(v.w, v.w, v.w, v.w);
    return add0 + add1;
}

// gtc/matrix_inverse.inl:95-147
inline M4 inverse_transpose(const M4& M) {
#define m(c_, r_) M.at(c_, r_)
    float S00 = m(2,2) * m(3,3) - m(3,2) * m(2,3);
    float S01 = m(2,1) * m(3,3) - m(3,1) * m(2,3);
    float S02 = m(2,1) * m(3,2) - m(3,1) * m(2,2);
    float S03 = m(2,0) * m(3,3) - m(3,0) * m(2,3);
    float S04 = m(2,0) * m(3,2) - m(3,0) * m(2,2);
    float S05 = m(2,0) * m(3,1) - m(3,0) * m(2,1);
    float S06 = m(1,2) * m(3,3) - m(3,2) * m(1,3);
    float S07 = m(1,1) * m(3,3) - m(3,1) * m(1,3);
    float S08 = m(1,1) * m(3,2) - m(3,1) * m(1,2);
    float S09 = m(1,0) * m(3,3) - m(3,0) * m(1,3);
    float S10 = m(1,0) * m(3,2) - m(3,0) * m(1,2);
    float S11 = m(1,1) * m(3,3) - m(3,1) * m(1,3);
    float S12 = m(1,0) * m(3,1) - m(3,0) * m(1,1);
    float S13 = m(1,2) * m(2,3) - m(2,2) * m(1,3);
    float S14 = m(1,1) * m(2,3) - m(2,1) * m(1,3);
    float S15 = m(1,1) * m(2,2) - m(2,1) * m(1,2);
    float S16 = m(1,0) * m(2,3) - m(2,0) * m(1,3);
    float S17 = m(1,0) * m(2,2) - m(2,0) * m(1,2);
    float S18 = m(1,0) * m(2,1) - m(2,0) * m(1,1);
    float I[4][4];
    I[0][0] = +(m(1,1) * S00 - m(1,2) * S01 + m(1,3) * S02);
    I[0][1] = -(m(1,0) * S00 - m(1,2) * S03 + m(1,3) * S04);
    I[0][2] = +(m(1,0) * S01 - m(1,1) * S03 + m(1,3) * S05);
    I[0][3] = -(m(1,0) * S02 - m(1,1) * S04 + m(1,2) * S05);
    I[1][0] = -(m(0,1) * S00 - m(0,2) * S01 + m(0,3) * S02);
    I[1][1] = +(m(0,0) * S00 - m(0,2) * S03 + m(0,3) * S04);
    I[1][2] = -(m(0,0) * S01 - m(0,1) * S03 + m(0,3) * S05);
    I[1][3] = +(m(0,0) * S02 - m(0,1) * S04 + m(0,2) * S05);
    I[2][0] = +(m(0,1) * S06 - m(0,2) * S07 + m(0,3) * S08);
    I[2][1] = -(m(0,0) * S06 - m(0,2) * S09 + m(0,3) * S10);
    I[2][2] = +(m(0,0) * S11 - m(0,1) * S09 + m(0,3) * S12);
    I[2][3] = -(m(0,0) * S08 - m(0,1) * S10 + m(0,2) * S12);
    I[3][0] = -(m(0,1) * S13 - m(0,2) * S14 + m(0,3) * S15);
    I[3][1] = +(m(0,0) * S13 - m(0,2) * S16 + m(0,3) * S17);
    I[3][2] = -(m(0,0) * S14 - m(0,1) * S16 + m(0,3) * S18);
    I[3][3] = +(m(0,0) * S15 - m(0,1) * S17 + m(0,2) * S18);
    float det = +m(0,0) * I[0][0] + m(0,1) * I[0][1] + m(0,2) * I[0][2] + m(0,3) * I[0][3];
#undef m
    M4 r;
    for (int i = 0; i < 4; i++) r.c[i] = V4(I[i][0] / det, I[i][1] / det, I[i][2] / det, I[i][3] / det);
    return r;
}

}  // namespace ptb
