// Linux stand-in for the part of Apple's <simd/simd.h> that the reference's host API exposes
// (simd_float4x4 in InstancedSplatRenderer::render, src/instanced_splat_renderer.h:21-26; the vector
// helpers TrackballCamera uses, src/trackball_camera.mm:6-8,30,34-36,65-84,137-162).
// Plain structs, column-major matrices (columns[c] = column c), every product accumulated left to
// right in fp32 so results are reproducible across compilers (build with -ffp-contract=off).
#pragma once
#include <cmath>

struct simd_float2 { float x, y; };
struct simd_float3 { float x, y, z; };
struct simd_float4 { float x, y, z, w; };
struct simd_float4x4 { simd_float4 columns[4]; };
struct simd_quatf { simd_float4 vector; };  // imaginary xyz, real w

inline simd_float2 simd_make_float2(float x, float y) { return {x, y}; }
inline simd_float3 simd_make_float3(float x, float y, float z) { return {x, y, z}; }
inline simd_float4 simd_make_float4(float x, float y, float z, float w) { return {x, y, z, w}; }

inline simd_float2 operator-(simd_float2 a, simd_float2 b) { return {a.x - b.x, a.y - b.y}; }
inline simd_float3 operator+(simd_float3 a, simd_float3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline simd_float3 operator-(simd_float3 a, simd_float3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline simd_float3 operator*(simd_float3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline simd_float3 operator*(float s, simd_float3 a) { return {s * a.x, s * a.y, s * a.z}; }
inline simd_float3& operator+=(simd_float3& a, simd_float3 b) { a = a + b; return a; }

inline float simd_dot(simd_float3 a, simd_float3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline float simd_length(simd_float3 a) { return std::sqrt(simd_dot(a, a)); }
inline simd_float3 simd_normalize(simd_float3 a) { const float l = simd_length(a); return {a.x / l, a.y / l, a.z / l}; }
inline simd_float3 simd_cross(simd_float3 a, simd_float3 b) {
    return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
inline float simd_clamp(float v, float lo, float hi) { return v < lo ? lo : (v > hi ? hi : v); }

inline simd_quatf simd_quaternion(float ix, float iy, float iz, float r) { return {{ix, iy, iz, r}}; }
inline simd_quatf simd_quaternion(float angle, simd_float3 axis) {
    const float s = std::sin(angle * 0.5f), c = std::cos(angle * 0.5f);
    return {{axis.x * s, axis.y * s, axis.z * s, c}};
}
inline simd_quatf simd_mul(simd_quatf p, simd_quatf q) {
    const simd_float4 a = p.vector, b = q.vector;
    return {{((a.w * b.x + a.x * b.w) + a.y * b.z) - a.z * b.y,
             ((a.w * b.y - a.x * b.z) + a.y * b.w) + a.z * b.x,
             ((a.w * b.z + a.x * b.y) - a.y * b.x) + a.z * b.w,
             ((a.w * b.w - a.x * b.x) - a.y * b.y) - a.z * b.z}};
}
inline simd_float3 simd_act(simd_quatf q, simd_float3 v) {   // v + 2 r (u x v) + 2 u x (u x v)
    const simd_float3 u = {q.vector.x, q.vector.y, q.vector.z};
    const simd_float3 t = simd_cross(u, v) * 2.0f;
    return (v + t * q.vector.w) + simd_cross(u, t);
}

inline simd_float4x4 simd_matrix(simd_float4 c0, simd_float4 c1, simd_float4 c2, simd_float4 c3) { return {{c0, c1, c2, c3}}; }
static const simd_float4x4 matrix_identity_float4x4 = {{{1, 0, 0, 0}, {0, 1, 0, 0}, {0, 0, 1, 0}, {0, 0, 0, 1}}};
inline simd_float4x4 simd_mul(const simd_float4x4& a, const simd_float4x4& b) {
    simd_float4x4 o;
    const float* A = &a.columns[0].x; const float* B = &b.columns[0].x; float* O = &o.columns[0].x;
    for (int j = 0; j < 4; ++j)
        for (int r = 0; r < 4; ++r)
            O[j * 4 + r] = ((A[r] * B[j * 4] + A[4 + r] * B[j * 4 + 1]) + A[8 + r] * B[j * 4 + 2]) + A[12 + r] * B[j * 4 + 3];
    return o;
}
