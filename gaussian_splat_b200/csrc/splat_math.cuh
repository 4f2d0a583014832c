// Pinned per-splat / per-fragment arithmetic.
// Compiled with -fmad=false: every + - * / sqrt below is one correctly rounded
// IEEE fp32 operation in the written order; fmaf() appears only where written.
// Follows shaders/gaussian_splat_tile.metal:40-157 (vertex_main and helpers),
// evaluated once per splat instead of once per quad corner.
#pragma once
#include "splat_internal.cuh"

// exp(-0.5*q), q in [0, ~9.3], as 2^t with t = q * (-0.5*log2 e): magic-add rounding (no cvt
// instructions), f = t - n exact, degree-5 minimax 2^f in explicit fmaf, exponent added as an
// integer. 10 instructions; max rel. error 9e-8 + 4e-7 from the rounding of t.
// c5 = the leading coefficient 0.0013264764565974474f (0x3AADDD2A), passed in so that a caller's hot loop can
// keep it in a register (an FFMA takes one immediate; the compiler otherwise re-materialises it per call).
#define EXP_NEG_HALF_C5 0.0013264764565974474f
__device__ __forceinline__ float exp_neg_half(float q, float c5 = EXP_NEG_HALF_C5) {
    const float t = q * -0.72134752044448170368f;
    const float s = t + 12582912.0f;
    const float n = s - 12582912.0f;
    const float f = t - n;
    float p = c5;
    p = fmaf(p, f, 0.009671512991189957f);
    p = fmaf(p, f, 0.05550733581185341f);
    p = fmaf(p, f, 0.24022242426872253f);
    p = fmaf(p, f, 0.6931470036506653f);
    p = fmaf(p, f, 1.0f);
    return __int_as_float(__float_as_int(p) + (__float_as_int(s) << 23));
}

__device__ __forceinline__ void mat4_point(const float* M, float x, float y, float z, float out[4]) {
#pragma unroll
    for (int r = 0; r < 4; ++r)
        out[r] = ((M[0 * 4 + r] * x + M[1 * 4 + r] * y) + M[2 * 4 + r] * z) + M[3 * 4 + r];
}

// View-dependent colour, SH degree 1..3 (build-side extension, the reference drops
// f_rest at src/instanced_splat_renderer.mm:367-370). Standard 3DGS real-SH basis on
// dir = normalize(pos - camera). b[k] = basis function k+1; every product/sum below is rounded
// separately in the written order (the CPU restatement carries the same text).
__device__ __forceinline__ void sh_basis(float x, float y, float z, int degree, float b[15]) {
    b[0] = -0.4886025119029199f * y;
    b[1] = 0.4886025119029199f * z;
    b[2] = -0.4886025119029199f * x;
    if (degree < 2) return;
    const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
    b[3] = 1.0925484305920792f * xy;
    b[4] = -1.0925484305920792f * yz;
    b[5] = 0.31539156525252005f * ((2.0f * zz - xx) - yy);
    b[6] = -1.0925484305920792f * xz;
    b[7] = 0.5462742152960396f * (xx - yy);
    if (degree < 3) return;
    b[8] = (-0.5900435899266435f * y) * (3.0f * xx - yy);
    b[9] = (2.890611442640554f * xy) * z;
    b[10] = (-0.4570457994644658f * y) * ((4.0f * zz - xx) - yy);
    b[11] = (0.3731763325901154f * z) * ((2.0f * zz - 3.0f * xx) - 3.0f * yy);
    b[12] = (-0.4570457994644658f * x) * ((4.0f * zz - xx) - yy);
    b[13] = (1.445305721320277f * z) * (xx - yy);
    b[14] = (-0.5900435899266435f * x) * (xx - 3.0f * yy);
}

// One splat -> 48-byte record. Returns tiles touched (0 = culled; rec untouched).
__device__ __forceinline__ unsigned int preprocess_one(
    float qw, float qx, float qy, float qz, float s0, float s1, float s2,
    float px, float py, float pz, unsigned int rg_half2, unsigned int b_half, float opacity,
    const FrameUniforms& U, splat_record& rec)
{
    // gaussian_splat_tile.metal:94-105
    float vp[4];
    mat4_point(U.V, px, py, pz, vp);
    const float zFront = -vp[2];
    if (!(zFront >= 1e-4f)) return 0;
    // the fragment stage discards in.depth < 0.001 (gaussian_splat_tile.metal:187, gaussian_splat_50layer.metal:154,
    // gaussian_splat.metal:209) and depth = zFront is constant over the quad: such a splat draws nothing in any mode
    if (zFront < 0.001f) return 0;

    // :40-49 quaternionToMatrix (normalize = v / sqrt(dot))
    const float qlen = sqrtf(((qw * qw + qx * qx) + qy * qy) + qz * qz);
    const float w = qw / qlen, x = qx / qlen, y = qy / qlen, z = qz / qlen;
    float R[3][3];
    R[0][0] = 1.0f - 2.0f * (y * y + z * z); R[1][0] = 2.0f * (x * y + w * z); R[2][0] = 2.0f * (x * z - w * y);
    R[0][1] = 2.0f * (x * y - w * z); R[1][1] = 1.0f - 2.0f * (x * x + z * z); R[2][1] = 2.0f * (y * z + w * x);
    R[0][2] = 2.0f * (x * z + w * y); R[1][2] = 2.0f * (y * z - w * x); R[2][2] = 1.0f - 2.0f * (x * x + y * y);

    // :51-60 M = R*S ; Sigma = M*M^T
    const float sc[3] = {s0, s1, s2};
    float M[3][3], Sg[3][3];
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) M[r][c] = R[r][c] * sc[c];
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) Sg[r][c] = (M[r][0] * M[c][0] + M[r][1] * M[c][1]) + M[r][2] * M[c][2];

    // :109-115 viewCov = (viewRot*Sigma)*viewRot^T ; Wm[r][c] = V[c][r]
    float Wm[3][3], T[3][3], C[3][3];
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) Wm[r][c] = U.V[c * 4 + r];
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) T[r][c] = (Wm[r][0] * Sg[0][c] + Wm[r][1] * Sg[1][c]) + Wm[r][2] * Sg[2][c];
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) C[r][c] = (T[r][0] * Wm[c][0] + T[r][1] * Wm[c][1]) + T[r][2] * Wm[c][2];

    // :117-123 Jacobian rows (literal sign kept)
    const float focalX = U.P[0] * (U.W * 0.5f);
    const float focalY = U.P[5] * (U.H * 0.5f);
    const float invZ = 1.0f / zFront;
    const float invZ2 = invZ * invZ;
    const float j00 = focalX * invZ, j02 = -((focalX * vp[0]) * invZ2);
    const float j11 = focalY * invZ, j12 = -((focalY * vp[1]) * invZ2);

    // :125-131 (exact-zero terms dropped)
    const float t0 = C[0][0] * j00 + C[0][2] * j02;
    const float t2 = C[2][0] * j00 + C[2][2] * j02;
    float a = j00 * t0 + j02 * t2;
    const float u0 = C[0][1] * j11 + C[0][2] * j12;
    const float u1 = C[1][1] * j11 + C[1][2] * j12;
    const float u2 = C[2][1] * j11 + C[2][2] * j12;
    const float b = j00 * u0 + j02 * u2;
    float c = j11 * u1 + j12 * u2;
    a = a + 1e-4f;
    c = c + 1e-4f;

    // :62-83 eigenSym2x2, :136-140
    const float tr = a + c;
    const float det = a * c - b * b;
    const float disc = fmaxf(0.0f, (0.25f * tr) * tr - det);
    const float s = sqrtf(disc);
    float l1 = 0.5f * tr + s;
    float l2 = 0.5f * tr - s;
    float e1x, e1y;
    if (fabsf(b) > 1e-8f) {
        const float ex = l1 - c, ey = b;
        const float el = sqrtf(ex * ex + ey * ey);
        e1x = ex / el; e1y = ey / el;
    } else if (a >= c) { e1x = 1.0f; e1y = 0.0f; }
    else { e1x = 0.0f; e1y = 1.0f; }
    l1 = fmaxf(l1, 0.0f);
    l2 = fmaxf(l2, 0.0f);
    if (!(l1 > 0.0f && l2 > 0.0f)) return 0;
    const float sd1 = sqrtf(l1), sd2 = sqrtf(l2);
    const float r1 = 3.0f * sd1, r2 = 3.0f * sd2;
    const float is1 = 1.0f / sd1, is2 = 1.0f / sd2;

    // :145-152 clip / NDC ; Metal clip volume 0 <= z <= w with w = 1 (D8)
    float cl[4];
    mat4_point(U.VP, px, py, pz, cl);
    const float invW = 1.0f / cl[3];
    const float ndcx = cl[0] * invW, ndcy = cl[1] * invW, ndcz = cl[2] * invW;
    if (!(ndcz >= 0.0f && ndcz <= 1.0f)) return 0;

    // viewport transform (y down) and tile rectangle (D9)
    const float cx = (ndcx * 0.5f + 0.5f) * U.W;
    const float cy = (0.5f - 0.5f * ndcy) * U.H;
    // D9': the quad's AABB (:142-143) intersected with the AABB of the ellipse u*u + v*v <= 2 ln 100 (:191), whose half
    // extents are (rho / 3) sqrt((r1 e1x)^2 + (r2 e2x)^2), rho / 3 = 1.011618 rounded UP to 1.0117
    const float ax1 = r1 * e1x, ax2 = r2 * e1y, ay1 = r1 * e1y, ay2 = r2 * e1x;
    const float hx = fminf(fabsf(ax1) + fabsf(ax2), 1.0117f * sqrtf(ax1 * ax1 + ax2 * ax2));
    const float hy = fminf(fabsf(ay1) + fabsf(ay2), 1.0117f * sqrtf(ay1 * ay1 + ay2 * ay2));
    float fi0 = ceilf((cx - hx) - 0.5f), fi1 = floorf((cx + hx) - 0.5f);
    float fj0 = ceilf((cy - hy) - 0.5f), fj1 = floorf((cy + hy) - 0.5f);
    const float Wm1 = U.W - 1.0f, Hm1 = U.H - 1.0f;
    if (!(fi1 >= 0.0f && fi0 <= Wm1 && fj1 >= 0.0f && fj0 <= Hm1 && fi1 >= fi0 && fj1 >= fj0)) return 0;
    fi0 = fmaxf(fi0, 0.0f); fi1 = fminf(fi1, Wm1);
    fj0 = fmaxf(fj0, 0.0f); fj1 = fminf(fj1, Hm1);
    int tx0 = (int)fi0 >> 4, tx1 = (int)fi1 >> 4, ty0 = (int)fj0 >> 4, ty1 = (int)fj1 >> 4;
    if (ty0 < U.row_begin) ty0 = U.row_begin;
    if (ty1 > U.row_end - 1) ty1 = U.row_end - 1;
    if (ty1 < ty0) return 0;

    rec.cx = cx; rec.cy = cy;
    rec.ux = e1x * is1;    rec.uy = -(e1y * is1);
    rec.vx = -(e1y * is2); rec.vy = -(e1x * is2);
    rec.opacity = opacity;
    rec.rg = rg_half2;
    rec.bd = b_half | ((unsigned int)__half_as_ushort(__float2half_rn(zFront)) << 16);
    rec.z_front = zFront;
    rec.rect_lo = (unsigned int)tx0 | ((unsigned int)ty0 << 16);
    rec.rect_hi = (unsigned int)tx1 | ((unsigned int)ty1 << 16);
    return (unsigned int)((tx1 - tx0 + 1) * (ty1 - ty0 + 1));
}

// D9' per-tile test: can the splat cover a pixel centre of tile (tx, ty)? The minimum of q = u*u + v*v over the tile's
// pixel-centre box lies on a box edge facing the splat centre (q grows along every ray from the centre), i.e. on the
// vertical line x = clamp(cx) or the horizontal line y = clamp(cy); on each line the minimiser is the clamped root of
// dq/dt = 0 (kx = B / C, ky = B / A of the quadratic form, computed once per splat by tile_ratios) and q is evaluated
// there with the fragment's own formula. The threshold carries a slack proportional to that formula's rounding error at
// this distance, so a tile holding a covered pixel is never dropped: images do not change, S2 stays exact per pixel.
// The CPU restatement carries the same text (tile_touch there too); the set of pairs is bit-identical.
__device__ __forceinline__ void tile_ratios(float ux, float uy, float vx, float vy, float& kx, float& ky) {
    const float A = ux * ux + vx * vx, B = ux * uy + vx * vy, C = uy * uy + vy * vy;
    kx = B / C; ky = B / A;
}
__device__ __forceinline__ bool tile_touch(float cx, float cy, float ux, float uy, float vx, float vy, float kx, float ky,
                                           int tx, int ty, int w, int h) {
    const float X0 = (float)(tx * 16) + 0.5f, X1 = (float)min(tx * 16 + 15, w - 1) + 0.5f;
    const float Y0 = (float)(ty * 16) + 0.5f, Y1 = (float)min(ty * 16 + 15, h - 1) + 0.5f;
    const float xc = fminf(fmaxf(cx, X0), X1), yc = fminf(fmaxf(cy, Y0), Y1);
    const float dxa = xc - cx;
    const float dya = fminf(fmaxf(-(dxa * kx), Y0 - cy), Y1 - cy);
    const float ua = fmaf(dxa, ux, dya * uy), va = fmaf(dxa, vx, dya * vy);
    const float qa = fmaf(ua, ua, va * va);
    const float dyb = yc - cy;
    const float dxb = fminf(fmaxf(-(dyb * ky), X0 - cx), X1 - cx);
    const float ub = fmaf(dxb, ux, dyb * uy), vb = fmaf(dxb, vx, dyb * vy);
    const float qb = fmaf(ub, ub, vb * vb);
    const float m = fmaxf(fabsf(cx - X0), fabsf(cx - X1)) + fmaxf(fabsf(cy - Y0), fabsf(cy - Y1));
    const float e = (m * (((fabsf(ux) + fabsf(uy)) + fabsf(vx)) + fabsf(vy))) * 9.5367431640625e-07f;   // 2^-20
    const float thr = (9.25f + 8.0f * e) + e * e;
    return !(fminf(qa, qb) > thr);
}
