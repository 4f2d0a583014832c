// Linux stand-in for the slice of Apple <simd/simd.h> that the reference's
// src/trackball_camera.{h,mm} uses, so that file compiles unmodified into oracle/_ref/
// (TEST INFRASTRUCTURE ONLY). Apple's library is a closed OS component: the
// choices below (normalize = v / sqrt(dot), products accumulated left to right) are ours.
#ifndef ORACLE_REF_SHIM_SIMD_H
#define ORACLE_REF_SHIM_SIMD_H
#include <math.h>

struct simd_float2 { float x, y; };
struct simd_float3 { float x, y, z; };
struct simd_float4 { float x, y, z, w; };
struct simd_float4x4 { simd_float4 columns[4]; };
struct simd_quatf { simd_float4 vector; };   // (ix, iy, iz, r)

static inline simd_float2 simd_make_float2(float x, float y) { return simd_float2{x, y}; }
static inline simd_float3 simd_make_float3(float x, float y, float z) { return simd_float3{x, y, z}; }
static inline simd_float4 simd_make_float4(float x, float y, float z, float w) { return simd_float4{x, y, z, w}; }

static inline simd_float2 operator-(simd_float2 a, simd_float2 b) { return simd_float2{a.x - b.x, a.y - b.y}; }
static inline simd_float3 operator-(simd_float3 a, simd_float3 b) { return simd_float3{a.x - b.x, a.y - b.y, a.z - b.z}; }
static inline simd_float3 operator+(simd_float3 a, simd_float3 b) { return simd_float3{a.x + b.x, a.y + b.y, a.z + b.z}; }
static inline simd_float3 operator*(simd_float3 a, float s) { return simd_float3{a.x * s, a.y * s, a.z * s}; }
static inline simd_float3 operator*(float s, simd_float3 a) { return simd_float3{s * a.x, s * a.y, s * a.z}; }
static inline simd_float3& operator+=(simd_float3& a, simd_float3 b) { a = a + b; return a; }

static inline float simd_dot(simd_float3 a, simd_float3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
static inline float simd_length(simd_float3 a) { return sqrtf(simd_dot(a, a)); }
static inline simd_float3 simd_normalize(simd_float3 a) { const float l = simd_length(a); return simd_float3{a.x / l, a.y / l, a.z / l}; }
static inline simd_float3 simd_cross(simd_float3 a, simd_float3 b) {
    return simd_float3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
static inline float simd_clamp(float x, float lo, float hi) { return x < lo ? lo : (x > hi ? hi : x); }

static inline simd_quatf simd_quaternion(float ix, float iy, float iz, float r) { return simd_quatf{simd_float4{ix, iy, iz, r}}; }
static inline simd_quatf simd_quaternion(float angle, simd_float3 axis) {
    const float s = sinf(angle * 0.5f), c = cosf(angle * 0.5f);
    return simd_quatf{simd_float4{axis.x * s, axis.y * s, axis.z * s, c}};
}
static inline simd_quatf simd_mul(simd_quatf p, simd_quatf q) {
    const simd_float4 a = p.vector, b = q.vector;
    return simd_quatf{simd_float4{
        ((a.w * b.x + a.x * b.w) + a.y * b.z) - a.z * b.y,
        ((a.w * b.y - a.x * b.z) + a.y * b.w) + a.z * b.x,
        ((a.w * b.z + a.x * b.y) - a.y * b.x) + a.z * b.w,
        ((a.w * b.w - a.x * b.x) - a.y * b.y) - a.z * b.z}};
}
// v' = v + 2 r (q x v) + 2 q x (q x v)
static inline simd_float3 simd_act(simd_quatf q, simd_float3 v) {
    const simd_float3 u = simd_float3{q.vector.x, q.vector.y, q.vector.z};
    const simd_float3 t = simd_cross(u, v) * 2.0f;
    return (v + t * q.vector.w) + simd_cross(u, t);
}

static inline simd_float4x4 simd_matrix(simd_float4 a, simd_float4 b, simd_float4 c, simd_float4 d) { return simd_float4x4{{a, b, c, d}}; }
static const simd_float4x4 matrix_identity_float4x4 = {{{1, 0, 0, 0}, {0, 1, 0, 0}, {0, 0, 1, 0}, {0, 0, 0, 1}}};
static inline simd_float4x4 simd_mul(simd_float4x4 a, simd_float4x4 b) {
    simd_float4x4 o;
    const float* A = &a.columns[0].x; const float* B = &b.columns[0].x; float* O = &o.columns[0].x;
    for (int j = 0; j < 4; ++j) for (int r = 0; r < 4; ++r)
        O[j * 4 + r] = ((A[0 * 4 + r] * B[j * 4 + 0] + A[1 * 4 + r] * B[j * 4 + 1]) + A[2 * 4 + r] * B[j * 4 + 2]) + A[3 * 4 + r] * B[j * 4 + 3];
    return o;
}
#endif
