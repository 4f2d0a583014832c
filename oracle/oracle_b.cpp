/* Oracle-B implementation. See oracle_b.h for scope, citations and the
 * "test infrastructure only / parity unpinned" notice.
 *
 * Build: g++ -O2 -std=c++17 -ffp-contract=off (no fast-math). Every fp32
 * operation below is a single correctly rounded IEEE operation in a fixed
 * order (D5); fmaf() is used only where written. Half arithmetic is done as
 * float-op-then-round, which is correctly rounded for binary16 because
 * 24 >= 2*11+2 (D7).
 */
#include "oracle_b.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <thread>
#include <vector>

namespace {

/* ------------------------------------------------------------------ half */
inline uint16_t f2h(float f) { _Float16 h = (_Float16)f; uint16_t b; std::memcpy(&b, &h, 2); return b; }
inline float h2f(uint16_t b) { _Float16 h; std::memcpy(&h, &b, 2); return (float)h; }
inline uint16_t hadd(uint16_t a, uint16_t b) { return f2h(h2f(a) + h2f(b)); }
inline uint16_t hsub(uint16_t a, uint16_t b) { return f2h(h2f(a) - h2f(b)); }
inline uint16_t hmul(uint16_t a, uint16_t b) { return f2h(h2f(a) * h2f(b)); }
const uint16_t H_ONE = 0x3C00;
/* 0.99h and 0.01h as the Metal compiler would round the literals. */
const uint16_t H_099 = 0x3BEC; /* 0.990234375 = RNE(0.99); SURVEY.md quotes the truncated 0x3BEB */
const uint16_t H_001 = 0x211F; /* 0.010002136   */

/* ------------------------------------------------------------ pinned exp */
/* exp(-0.5*q) for q in [0, ~9.3] as 2^t, t = q * (-0.5*log2 e): round-to-int by the
 * 1.5*2^23 magic add (no cvt), f = t - n in [-0.5, 0.5] exact, degree-5 minimax 2^f in
 * explicit fmaf (max rel. error 9e-8), exponent added as an integer. The CUDA kernel
 * carries the same constants and sequence (splat_math.cuh). */
inline float exp_neg_half(float q) {
    const float t = q * -0.72134752044448170368f;
    const float s = t + 12582912.0f;
    const float n = s - 12582912.0f;
    const float f = t - n;
    float p = 0.0013264764565974474f;
    p = std::fmaf(p, f, 0.009671512991189957f);
    p = std::fmaf(p, f, 0.05550733581185341f);
    p = std::fmaf(p, f, 0.24022242426872253f);
    p = std::fmaf(p, f, 0.6931470036506653f);
    p = std::fmaf(p, f, 1.0f);
    uint32_t pb, sb;
    std::memcpy(&pb, &p, 4); std::memcpy(&sb, &s, 4);
    pb += sb << 23;                       /* low bits of s hold n (two's complement) */
    float r; std::memcpy(&r, &pb, 4);
    return r;
}

/* ------------------------------------------------------------ S1 helpers */
struct Uni {
    float V[16], P[16], VP[16];
    float W, H;
    int w, h, tiles_x, tiles_y, row_begin, row_end;
    float cam[3];
    int tight;   /* 1 = D9' (ellipse-clipped rectangle + per-tile test), 0 = the plain quad AABB of D9 */
};

/* M * (x,y,z,1), column-major, k = 0,1,2,3 left to right (D5). */
inline void mat4_point(const float* M, float x, float y, float z, float out[4]) {
    for (int r = 0; r < 4; ++r)
        out[r] = ((M[0 * 4 + r] * x + M[1 * 4 + r] * y) + M[2 * 4 + r] * z) + M[3 * 4 + r];
}

/* D11 extension: real-SH basis functions 1..15 on a unit direction (standard 3DGS constants);
 * same text as gaussian_splat_b200/csrc/splat_math.cuh: sh_basis. */
void sh_basis(float x, float y, float z, int degree, float b[15]) {
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

/* colour = max(0, loader colour + sum_k basis_k(dir) * f_rest[channel*15 + k]), accumulated in
 * coefficient-major order (k outer, channel inner), then rounded to half. */
void sh_colour(const float* inst, const float* sh45, const float cam[3], int degree, orb_record& rec) {
    const float dx = inst[7] - cam[0], dy = inst[8] - cam[1], dz = inst[9] - cam[2];
    const float dl = std::sqrt((dx * dx + dy * dy) + dz * dz);
    float bs[15];
    sh_basis(dx / dl, dy / dl, dz / dl, degree, bs);
    float col[3] = {inst[10], inst[11], inst[12]};
    const int n_coef = (degree + 1) * (degree + 1) - 1;
    for (int k = 0; k < n_coef; ++k)
        for (int c = 0; c < 3; ++c) col[c] = col[c] + bs[k] * sh45[c * 15 + k];
    rec.rg = (uint32_t)f2h(std::fmax(col[0], 0.0f)) | ((uint32_t)f2h(std::fmax(col[1], 0.0f)) << 16);
    rec.bd = (rec.bd & 0xFFFF0000u) | (uint32_t)f2h(std::fmax(col[2], 0.0f));
}

/* One splat. Returns tiles touched (0 = culled) and fills rec. */
uint32_t preprocess_one(const float* inst, const Uni& U, orb_record& rec) {
    const float qw = inst[0], qx = inst[1], qy = inst[2], qz = inst[3];
    const float s0 = inst[4], s1 = inst[5], s2 = inst[6];
    const float px = inst[7], py = inst[8], pz = inst[9];

    /* gaussian_splat_tile.metal:94-105 */
    float vp[4];
    mat4_point(U.V, px, py, pz, vp);
    const float zFront = -vp[2];
    if (!(zFront >= 1e-4f)) return 0;
    /* the fragment stage discards in.depth < 0.001 (gaussian_splat_tile.metal:187, gaussian_splat_50layer.metal:154,
     * gaussian_splat.metal:209) and depth = zFront is constant over the quad: such a splat draws nothing in any mode */
    if (zFront < 0.001f) return 0;

    /* :40-49 quaternionToMatrix, normalize = v / sqrt(dot) */
    const float qlen = std::sqrt(((qw * qw + qx * qx) + qy * qy) + qz * qz);
    const float w = qw / qlen, x = qx / qlen, y = qy / qlen, z = qz / qlen;
    float R[3][3]; /* R[row][col] */
    R[0][0] = 1.0f - 2.0f * (y * y + z * z); R[1][0] = 2.0f * (x * y + w * z); R[2][0] = 2.0f * (x * z - w * y);
    R[0][1] = 2.0f * (x * y - w * z); R[1][1] = 1.0f - 2.0f * (x * x + z * z); R[2][1] = 2.0f * (y * z + w * x);
    R[0][2] = 2.0f * (x * z + w * y); R[1][2] = 2.0f * (y * z - w * x); R[2][2] = 1.0f - 2.0f * (x * x + y * y);

    /* :51-60 M = R*S, Sigma = M*M^T */
    const float sc[3] = {s0, s1, s2};
    float M[3][3];
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) M[r][c] = R[r][c] * sc[c];
    float Sg[3][3];
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c)
        Sg[r][c] = (M[r][0] * M[c][0] + M[r][1] * M[c][1]) + M[r][2] * M[c][2];

    /* :109-115 viewCov = (viewRot*Sigma)*viewRot^T ; Wm[r][c] = V[c][r] */
    float Wm[3][3];
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) Wm[r][c] = U.V[c * 4 + r];
    float T[3][3], C[3][3];
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c)
        T[r][c] = (Wm[r][0] * Sg[0][c] + Wm[r][1] * Sg[1][c]) + Wm[r][2] * Sg[2][c];
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c)
        C[r][c] = (T[r][0] * Wm[c][0] + T[r][1] * Wm[c][1]) + T[r][2] * Wm[c][2];

    /* :117-123 Jacobian rows (literal sign quirk kept) */
    const float focalX = U.P[0] * (U.W * 0.5f);
    const float focalY = U.P[5] * (U.H * 0.5f);
    const float invZ = 1.0f / zFront;
    const float invZ2 = invZ * invZ;
    const float j00 = focalX * invZ, j02 = -((focalX * vp[0]) * invZ2);
    const float j11 = focalY * invZ, j12 = -((focalY * vp[1]) * invZ2);

    /* :125-131 (zero terms dropped: exact) */
    float t0 = C[0][0] * j00 + C[0][2] * j02;
    float t2 = C[2][0] * j00 + C[2][2] * j02;
    float a = j00 * t0 + j02 * t2;
    float u0 = C[0][1] * j11 + C[0][2] * j12;
    float u1 = C[1][1] * j11 + C[1][2] * j12;
    float u2 = C[2][1] * j11 + C[2][2] * j12;
    float b = j00 * u0 + j02 * u2;
    float c = j11 * u1 + j12 * u2;
    a = a + 1e-4f;
    c = c + 1e-4f;

    /* :62-83 eigenSym2x2, :136-140 */
    const float tr = a + c;
    const float det = a * c - b * b;
    const float disc = std::fmax(0.0f, (0.25f * tr) * tr - det);
    const float s = std::sqrt(disc);
    float l1 = 0.5f * tr + s;
    float l2 = 0.5f * tr - s;
    float e1x, e1y;
    if (std::fabs(b) > 1e-8f) {
        const float ex = l1 - c, ey = b;
        const float el = std::sqrt(ex * ex + ey * ey);
        e1x = ex / el; e1y = ey / el;
    } else if (a >= c) { e1x = 1.0f; e1y = 0.0f; }
    else { e1x = 0.0f; e1y = 1.0f; }
    l1 = std::fmax(l1, 0.0f);
    l2 = std::fmax(l2, 0.0f);
    if (!(l1 > 0.0f && l2 > 0.0f)) return 0; /* zero-area quad: no fragments */
    const float sd1 = std::sqrt(l1), sd2 = std::sqrt(l2);
    const float r1 = 3.0f * sd1, r2 = 3.0f * sd2;
    const float is1 = 1.0f / sd1, is2 = 1.0f / sd2;

    /* :145-152 clip / NDC ; D8 z clip (Metal clip volume 0 <= z <= w, w = 1) */
    float cl[4];
    mat4_point(U.VP, px, py, pz, cl);
    const float invW = 1.0f / cl[3];
    const float ndcx = cl[0] * invW, ndcy = cl[1] * invW, ndcz = cl[2] * invW;
    if (!(ndcz >= 0.0f && ndcz <= 1.0f)) return 0;

    /* viewport transform, y down (D9) */
    const float cx = (ndcx * 0.5f + 0.5f) * U.W;
    const float cy = (0.5f - 0.5f * ndcy) * U.H;
    /* D9': AABB of the quad (:142-143) intersected with the AABB of the ellipse u*u + v*v <= 2 ln 100 (:191):
     * half extents (rho/3) sqrt((r1 e1x)^2 + (r2 e2x)^2), rho/3 = 1.011618 rounded UP to 1.0117 */
    const float ax1 = r1 * e1x, ax2 = r2 * e1y, ay1 = r1 * e1y, ay2 = r2 * e1x;
    float hx = std::fabs(ax1) + std::fabs(ax2), hy = std::fabs(ay1) + std::fabs(ay2);
    if (U.tight) {
        hx = std::fmin(hx, 1.0117f * std::sqrt(ax1 * ax1 + ax2 * ax2));
        hy = std::fmin(hy, 1.0117f * std::sqrt(ay1 * ay1 + ay2 * ay2));
    }
    float fi0 = std::ceil((cx - hx) - 0.5f), fi1 = std::floor((cx + hx) - 0.5f);
    float fj0 = std::ceil((cy - hy) - 0.5f), fj1 = std::floor((cy + hy) - 0.5f);
    const float Wm1 = U.W - 1.0f, Hm1 = U.H - 1.0f;
    if (!(fi1 >= 0.0f && fi0 <= Wm1 && fj1 >= 0.0f && fj0 <= Hm1 && fi1 >= fi0 && fj1 >= fj0)) return 0;
    fi0 = std::fmax(fi0, 0.0f); fi1 = std::fmin(fi1, Wm1);
    fj0 = std::fmax(fj0, 0.0f); fj1 = std::fmin(fj1, Hm1);
    int tx0 = (int)fi0 >> 4, tx1 = (int)fi1 >> 4, ty0 = (int)fj0 >> 4, ty1 = (int)fj1 >> 4;
    /* band restriction (multi-GPU tile-row bands) */
    if (ty0 < U.row_begin) ty0 = U.row_begin;
    if (ty1 > U.row_end - 1) ty1 = U.row_end - 1;
    if (ty1 < ty0) return 0;

    rec.cx = cx; rec.cy = cy;
    rec.ux = e1x * is1;    rec.uy = -(e1y * is1);
    rec.vx = -(e1y * is2); rec.vy = -(e1x * is2);
    rec.opacity = inst[13];
    rec.rg = (uint32_t)f2h(inst[10]) | ((uint32_t)f2h(inst[11]) << 16);
    rec.bd = (uint32_t)f2h(inst[12]) | ((uint32_t)f2h(zFront) << 16);
    rec.z_front = zFront;
    rec.rect_lo = (uint32_t)tx0 | ((uint32_t)ty0 << 16);
    rec.rect_hi = (uint32_t)tx1 | ((uint32_t)ty1 << 16);
    return (uint32_t)((tx1 - tx0 + 1) * (ty1 - ty0 + 1));
}

/* D9' per-tile test: can the splat cover a pixel centre of tile (tx, ty)? Conservative: the minimum of
 * q = u*u + v*v over the tile's pixel-centre box is attained on a box edge facing the splat centre (q grows
 * along every ray from the centre), i.e. on the vertical line x = clamp(cx) or the horizontal line
 * y = clamp(cy); on each line the minimiser is the clamped root of dq/dt = 0, and q is evaluated there with the
 * fragment's own formula (:178-191). The threshold carries a slack proportional to the rounding error of that
 * formula at this distance, so a tile holding a covered pixel is never dropped (images do not change: S2 stays
 * exact per pixel). Same text as gaussian_splat_b200/csrc/splat_math.cuh: tile_touch. */
inline void tile_ratios(const orb_record& r, float& kx, float& ky) {
    const float A = r.ux * r.ux + r.vx * r.vx, B = r.ux * r.uy + r.vx * r.vy, C = r.uy * r.uy + r.vy * r.vy;
    kx = B / C; ky = B / A;
}
inline bool tile_touch(const orb_record& r, float kx, float ky, int tx, int ty, int w, int h) {
    const float X0 = (float)(tx * 16) + 0.5f, X1 = (float)std::min(tx * 16 + 15, w - 1) + 0.5f;
    const float Y0 = (float)(ty * 16) + 0.5f, Y1 = (float)std::min(ty * 16 + 15, h - 1) + 0.5f;
    const float xc = std::fmin(std::fmax(r.cx, X0), X1), yc = std::fmin(std::fmax(r.cy, Y0), Y1);
    const float dxa = xc - r.cx;
    const float dya = std::fmin(std::fmax(-(dxa * kx), Y0 - r.cy), Y1 - r.cy);
    const float ua = std::fmaf(dxa, r.ux, dya * r.uy), va = std::fmaf(dxa, r.vx, dya * r.vy);
    const float qa = std::fmaf(ua, ua, va * va);
    const float dyb = yc - r.cy;
    const float dxb = std::fmin(std::fmax(-(dyb * ky), X0 - r.cx), X1 - r.cx);
    const float ub = std::fmaf(dxb, r.ux, dyb * r.uy), vb = std::fmaf(dxb, r.vx, dyb * r.vy);
    const float qb = std::fmaf(ub, ub, vb * vb);
    const float m = std::fmax(std::fabs(r.cx - X0), std::fabs(r.cx - X1)) + std::fmax(std::fabs(r.cy - Y0), std::fabs(r.cy - Y1));
    const float e = (m * (((std::fabs(r.ux) + std::fabs(r.uy)) + std::fabs(r.vx)) + std::fabs(r.vy))) * 9.5367431640625e-07f;   /* 2^-20 */
    const float thr = (9.25f + 8.0f * e) + e * e;
    return !(std::fmin(qa, qb) > thr);
}

} // namespace

/* ================================================================ frame */
struct orb_frame {
    Uni U;
    orb_opts opts;
    uint32_t n = 0, T = 0;
    std::vector<orb_record> rec;
    std::vector<uint32_t> touched, offsets;
    std::vector<uint64_t> keys_u, keys_s;
    std::vector<uint32_t> vals_u, vals_s;
    std::vector<uint32_t> ranges;
};

namespace {

/* D3: depth field of the 64-bit key. */
inline uint32_t depth_field(const orb_opts& o, const orb_record& r, uint32_t idx) {
    if (o.mode == ORB_MODE_LAYER50 || o.mode == ORB_MODE_MLAB) return idx;   /* arrival = instance order */
    const uint32_t h = r.bd >> 16;
    return o.order == ORB_ORDER_REFERENCE ? (0xFFFFu - h) : h;
}

void lsd_sort(std::vector<uint64_t>& k, std::vector<uint32_t>& v) {
    /* stable by construction: equivalent to std::stable_sort on the key */
    const size_t n = k.size();
    std::vector<uint64_t> k2(n); std::vector<uint32_t> v2(n);
    for (int shift = 0; shift < 64; shift += 8) {
        size_t hist[257] = {0};
        for (size_t i = 0; i < n; ++i) hist[((k[i] >> shift) & 0xFF) + 1]++;
        if (hist[((n ? k[0] : 0) >> shift & 0xFF) + 1] == n) continue; /* constant digit */
        for (int d = 0; d < 256; ++d) hist[d + 1] += hist[d];
        for (size_t i = 0; i < n; ++i) { size_t p = hist[(k[i] >> shift) & 0xFF]++; k2[p] = k[i]; v2[p] = v[i]; }
        k.swap(k2); v.swap(v2);
    }
}

/* Per-tile blend, TILE mode (gaussian_splat_tile.metal:178-206 + :215-267)
 * under D1 (per-pixel lists) and D4 (cap K in sorted order). */
uint64_t blend_rows(const orb_frame& F, uint16_t* fb, int ty_begin, int ty_end) {
    const Uni& U = F.U;
    const int K = F.opts.max_fragments;
    const bool layer = F.opts.mode == ORB_MODE_LAYER50;
    const bool ref_order = F.opts.order == ORB_ORDER_REFERENCE;
    uint64_t tests = 0;
    struct Frag { uint32_t key; uint16_t r, g, b, a; };
    std::vector<Frag> list((size_t)K);
    for (int ty = ty_begin; ty < ty_end; ++ty)
    for (int tx = 0; tx < U.tiles_x; ++tx) {
        const uint32_t tile = (uint32_t)(ty * U.tiles_x + tx);
        const uint32_t beg = F.ranges[2 * tile], end = F.ranges[2 * tile + 1];
        for (int ly = 0; ly < 16; ++ly) for (int lx = 0; lx < 16; ++lx) {
            const int i = tx * 16 + lx, j = ty * 16 + ly;
            if (i >= U.w || j >= U.h) continue;
            const float pxc = (float)i + 0.5f, pyc = (float)j + 0.5f;
            uint16_t aR = 0, aG = 0, aB = 0;
            uint16_t acc = layer ? H_ONE : 0;  /* accumVis / accumAlpha */
            int cnt = 0;
            for (uint32_t e = beg; e < end; ++e) {
                const orb_record& r = F.rec[F.vals_s[e]];
                ++tests;
                const float dx = pxc - r.cx, dy = pyc - r.cy;
                const float u = std::fmaf(dx, r.ux, dy * r.uy);
                const float v = std::fmaf(dx, r.vx, dy * r.vy);
                const float q = std::fmaf(u, u, v * v);
                if (!(std::fabs(u) <= 3.0f && std::fabs(v) <= 3.0f && q <= 9.2103404f)) continue;
                const float alpha = exp_neg_half(q) * r.opacity;
                const uint16_t cr = (uint16_t)(r.rg & 0xFFFF), cg = (uint16_t)(r.rg >> 16), cb = (uint16_t)(r.bd & 0xFFFF);
                if (!layer) {
                    /* :252-264 */
                    const uint16_t ah = f2h(alpha);
                    const uint16_t src = hmul(ah, hsub(H_ONE, acc));
                    aR = hadd(aR, hmul(cr, src)); aG = hadd(aG, hmul(cg, src)); aB = hadd(aB, hmul(cb, src));
                    acc = hadd(acc, src);
                    ++cnt;
                    if (h2f(acc) >= h2f(H_099) || cnt >= K) break;
                } else {
                    /* 50layer :170-176: first K arrivals (instance order) kept */
                    const uint32_t h = r.bd >> 16;
                    const uint32_t dk = ref_order ? (0xFFFFu - h) : h;
                    list[cnt].key = (dk << 8) | (uint32_t)cnt;
                    list[cnt].r = cr; list[cnt].g = cg; list[cnt].b = cb;
                    list[cnt].a = f2h(1.0f - alpha);
                    if (++cnt >= K) break;
                }
            }
            uint16_t* out = fb + ((size_t)j * U.w + i) * 4;
            if (!layer) {
                out[0] = aR; out[1] = aG; out[2] = aB; out[3] = acc;   /* empty -> zeros */
            } else if (cnt == 0) {
                out[0] = out[1] = out[2] = out[3] = 0;                 /* :192-195 */
            } else {
                /* :198-206 stable insertion sort == ascending sort of unique (depthkey,slot) */
                std::sort(list.begin(), list.begin() + cnt, [](const Frag& a, const Frag& b) { return a.key < b.key; });
                for (int k = 0; k < cnt; ++k) {                         /* :209-220 */
                    aR = hadd(aR, hmul(list[k].r, acc)); aG = hadd(aG, hmul(list[k].g, acc)); aB = hadd(aB, hmul(list[k].b, acc));
                    acc = hmul(acc, list[k].a);
                    if (h2f(acc) < h2f(H_001)) break;
                }
                out[0] = aR; out[1] = aG; out[2] = aB; out[3] = hsub(H_ONE, acc);
            }
        }
    }
    return tests;
}

/* MLAB mode: streaming 6-layer K-buffer in instance order, then the resolve pass
 * (shaders/gaussian_splat.metal:201-291 fragment_main, :329-361 resolve_main; the eight RGBA16F attachments
 * are cleared to (0,0,0,1), src/instanced_splat_renderer.mm:536-541, so the depth slot of layer 3 - the
 * alpha channel of depths01 - starts at 1.0h and the others at 0). The framebuffer receives resolve_main's
 * return value; the fixed-function SrcAlpha blend over the drawable (:182-185) is presentation, not
 * part of this path. Fragment coverage is the pinned rule of the other two modes (D6). */
uint64_t blend_rows_mlab(const orb_frame& F, uint16_t* fb, int ty_begin, int ty_end) {
    const Uni& U = F.U;
    uint64_t tests = 0;
    struct Layer { uint16_t r, g, b, a; };
    for (int ty = ty_begin; ty < ty_end; ++ty)
    for (int tx = 0; tx < U.tiles_x; ++tx) {
        const uint32_t tile = (uint32_t)(ty * U.tiles_x + tx);
        const uint32_t beg = F.ranges[2 * tile], end = F.ranges[2 * tile + 1];
        for (int ly = 0; ly < 16; ++ly) for (int lx = 0; lx < 16; ++lx) {
            const int i = tx * 16 + lx, j = ty * 16 + ly;
            if (i >= U.w || j >= U.h) continue;
            const float pxc = (float)i + 0.5f, pyc = (float)j + 0.5f;
            Layer L[6]; uint16_t Dp[6];
            for (int k = 0; k < 6; ++k) { L[k] = Layer{0, 0, 0, H_ONE}; Dp[k] = 0; }
            Dp[3] = H_ONE;                                              /* clear colour alpha of depths01 */
            for (uint32_t e = beg; e < end; ++e) {
                const orb_record& r = F.rec[F.vals_s[e]];
                ++tests;
                if (r.z_front < 0.001f) continue;                       /* :209-211 */
                const float dx = pxc - r.cx, dy = pyc - r.cy;
                const float u = std::fmaf(dx, r.ux, dy * r.uy);
                const float v = std::fmaf(dx, r.vx, dy * r.vy);
                const float q = std::fmaf(u, u, v * v);
                if (!(std::fabs(u) <= 3.0f && std::fabs(v) <= 3.0f && q <= 9.2103404f)) continue;   /* :214-219 */
                const float alpha = exp_neg_half(q) * r.opacity;        /* :222 */
                const uint16_t ah = f2h(alpha);
                Layer nl{hmul((uint16_t)(r.rg & 0xFFFF), ah), hmul((uint16_t)(r.rg >> 16), ah),
                         hmul((uint16_t)(r.bd & 0xFFFF), ah), hsub(H_ONE, ah)};                     /* :226 */
                uint16_t nd = (uint16_t)(r.bd >> 16);                   /* half(in.depth) :223 */
                for (int k = 0; k < 6; ++k)                             /* :247-262 */
                    if (h2f(nd) >= h2f(Dp[k])) { std::swap(L[k], nl); std::swap(Dp[k], nd); }
                const bool closer = h2f(nd) >= h2f(Dp[5]);              /* :265-276 */
                const Layer front = closer ? nl : L[5], back = closer ? L[5] : nl;
                L[5] = Layer{hadd(back.r, hmul(front.r, back.a)), hadd(back.g, hmul(front.g, back.a)),
                             hadd(back.b, hmul(front.b, back.a)), hmul(front.a, back.a)};
                if (closer) Dp[5] = nd;
            }
            uint16_t cR = 0, cG = 0, cB = 0, at = H_ONE;                /* resolve_main :331-360 */
            for (int k = 0; k < 6; ++k) {
                cR = hadd(cR, hmul(L[k].r, at)); cG = hadd(cG, hmul(L[k].g, at)); cB = hadd(cB, hmul(L[k].b, at));
                at = hmul(at, L[k].a);
            }
            uint16_t* out = fb + ((size_t)j * U.w + i) * 4;
            out[0] = cR; out[1] = cG; out[2] = cB; out[3] = hsub(H_ONE, at);
        }
    }
    return tests;
}

} // namespace

static int g_threads = 1;

extern "C" {

void orb_set_threads(int n) { g_threads = n < 1 ? 1 : n; }

orb_frame* orb_frame_create(const float* inst, uint32_t n, const float* sh_rest,
                            const float* view, const float* proj, int w, int h,
                            const orb_opts* opts) {
    orb_frame* F = new orb_frame();
    F->opts = *opts;
    Uni& U = F->U;
    std::memcpy(U.V, view, 64); std::memcpy(U.P, proj, 64);
    orb_mat4_mul(proj, view, U.VP);                 /* instanced_splat_renderer.mm:453 */
    U.w = w; U.h = h; U.W = (float)w; U.H = (float)h;
    U.tiles_x = (w + 15) / 16; U.tiles_y = (h + 15) / 16;
    U.row_begin = 0; U.row_end = U.tiles_y;
    if (opts->tile_row_end > 0) { U.row_begin = opts->tile_row_begin; U.row_end = std::min(opts->tile_row_end, U.tiles_y); }
    U.tight = opts->tile_test;
    F->n = n; F->T = (uint32_t)(U.tiles_x * U.tiles_y);
    for (int j = 0; j < 3; ++j) {                   /* camera position = -R^T t */
        const float m0 = view[j * 4 + 0] * view[12], m1 = view[j * 4 + 1] * view[13], m2 = view[j * 4 + 2] * view[14];
        float acc = m0 + m1; acc = acc + m2;
        U.cam[j] = -acc;
    }
    const int sh_degree = (sh_rest && opts->sh_degree > 0) ? opts->sh_degree : 0;

    /* K1 */
    F->rec.assign(n, orb_record{}); F->touched.assign(n, 0); F->offsets.assign(n, 0);
    {
        auto work = [&](uint32_t b, uint32_t e) {
            for (uint32_t i = b; i < e; ++i) {
                orb_record r{};
                uint32_t t = preprocess_one(inst + (size_t)i * 14, U, r);
                if (t && sh_degree) sh_colour(inst + (size_t)i * 14, sh_rest + (size_t)i * 45, U.cam, sh_degree, r);
                if (t) F->rec[i] = r;
                F->touched[i] = t;
            }
        };
        const int nt = g_threads;
        if (nt <= 1 || n < 4096) work(0, n);
        else {
            std::vector<std::thread> th;
            for (int t = 0; t < nt; ++t)
                th.emplace_back(work, (uint32_t)((uint64_t)n * t / nt), (uint32_t)((uint64_t)n * (t + 1) / nt));
            for (auto& x : th) x.join();
        }
    }
    /* K2: duplication, row-major (ty outer, tx inner), of the tiles that pass the D9' test; touched[] becomes the
     * number of pairs emitted (a splat whose tiles all fail is culled: record zeroed), offsets[] their exclusive scan */
    {
        auto count = [&](uint32_t b, uint32_t e) {
            for (uint32_t i = b; i < e; ++i) {
                if (!F->touched[i] || !F->opts.tile_test) continue;     /* D9' off: every tile of the rectangle */
                const orb_record& r = F->rec[i];
                const uint32_t tx0 = r.rect_lo & 0xFFFF, ty0 = r.rect_lo >> 16, tx1 = r.rect_hi & 0xFFFF, ty1 = r.rect_hi >> 16;
                float kx, ky; tile_ratios(r, kx, ky);
                uint32_t c = 0;
                for (uint32_t ty = ty0; ty <= ty1; ++ty) for (uint32_t tx = tx0; tx <= tx1; ++tx)
                    c += tile_touch(r, kx, ky, (int)tx, (int)ty, U.w, U.h) ? 1u : 0u;
                F->touched[i] = c;
                if (!c) F->rec[i] = orb_record{};
            }
        };
        const int nt = g_threads;
        if (nt <= 1 || n < 4096) count(0, n);
        else {
            std::vector<std::thread> th;
            for (int t = 0; t < nt; ++t)
                th.emplace_back(count, (uint32_t)((uint64_t)n * t / nt), (uint32_t)((uint64_t)n * (t + 1) / nt));
            for (auto& x : th) x.join();
        }
    }
    uint64_t D = 0;
    for (uint32_t i = 0; i < n; ++i) { F->offsets[i] = (uint32_t)D; D += F->touched[i]; }
    F->keys_u.resize(D); F->vals_u.resize(D);
    for (uint32_t i = 0; i < n; ++i) {
        if (!F->touched[i]) continue;
        const orb_record& r = F->rec[i];
        const uint32_t tx0 = r.rect_lo & 0xFFFF, ty0 = r.rect_lo >> 16, tx1 = r.rect_hi & 0xFFFF, ty1 = r.rect_hi >> 16;
        const uint32_t df = depth_field(F->opts, r, i);
        float kx, ky; tile_ratios(r, kx, ky);
        uint64_t o = F->offsets[i];
        for (uint32_t ty = ty0; ty <= ty1; ++ty) for (uint32_t tx = tx0; tx <= tx1; ++tx) {
            if (F->opts.tile_test && !tile_touch(r, kx, ky, (int)tx, (int)ty, U.w, U.h)) continue;
            F->keys_u[o] = ((uint64_t)(ty * (uint32_t)U.tiles_x + tx) << 32) | df;
            F->vals_u[o] = i; ++o;
        }
    }
    /* K3 */
    F->keys_s = F->keys_u; F->vals_s = F->vals_u;
    lsd_sort(F->keys_s, F->vals_s);
    /* K4 */
    F->ranges.assign((size_t)F->T * 2, 0);
    for (uint64_t k = 0; k < D; ++k) {
        const uint32_t t = (uint32_t)(F->keys_s[k] >> 32);
        if (k == 0 || t != (uint32_t)(F->keys_s[k - 1] >> 32)) F->ranges[2 * t] = (uint32_t)k;
        if (k + 1 == D || t != (uint32_t)(F->keys_s[k + 1] >> 32)) F->ranges[2 * t + 1] = (uint32_t)(k + 1);
    }
    return F;
}

void orb_frame_destroy(orb_frame* F) { delete F; }
uint32_t orb_n(const orb_frame* F) { return F->n; }
uint64_t orb_num_pairs(const orb_frame* F) { return F->keys_u.size(); }
uint32_t orb_num_tiles(const orb_frame* F) { return F->T; }
const orb_record* orb_records(const orb_frame* F) { return F->rec.data(); }
const uint32_t* orb_tiles_touched(const orb_frame* F) { return F->touched.data(); }
const uint32_t* orb_offsets(const orb_frame* F) { return F->offsets.data(); }
const uint64_t* orb_keys_unsorted(const orb_frame* F) { return F->keys_u.data(); }
const uint32_t* orb_vals_unsorted(const orb_frame* F) { return F->vals_u.data(); }
const uint64_t* orb_keys_sorted(const orb_frame* F) { return F->keys_s.data(); }
const uint32_t* orb_vals_sorted(const orb_frame* F) { return F->vals_s.data(); }
const uint32_t* orb_tile_ranges(const orb_frame* F) { return F->ranges.data(); }

uint64_t orb_blend_rows(const orb_frame* F, uint16_t* fb, int row_begin, int row_end, int n_threads) {
    const int rb = std::max(row_begin, F->U.row_begin), re = std::min(row_end, F->U.row_end);
    auto blend_rows = [](const orb_frame& f, uint16_t* o, int a, int b) {
        return f.opts.mode == ORB_MODE_MLAB ? blend_rows_mlab(f, o, a, b) : ::blend_rows(f, o, a, b);
    };
    if (re <= rb) return 0;
    if (n_threads <= 1) return blend_rows(*F, fb, rb, re);
    std::vector<std::thread> th; std::vector<uint64_t> cnt((size_t)n_threads, 0);
    /* interleave rows for balance */
    for (int t = 0; t < n_threads; ++t)
        th.emplace_back([&, t] { for (int r = rb + t; r < re; r += n_threads) cnt[t] += blend_rows(*F, fb, r, r + 1); });
    uint64_t tot = 0;
    for (int t = 0; t < n_threads; ++t) { th[t].join(); tot += cnt[t]; }
    return tot;
}

uint64_t orb_blend(const orb_frame* F, uint16_t* fb, int n_threads) {
    return orb_blend_rows(F, fb, F->U.row_begin, F->U.row_end, n_threads);
}

/* src/trackball_camera.mm:136-148 */
void orb_look_at(const float eye[3], const float center[3], const float up[3], float o[16]) {
    auto norm = [](float v[3]) { float l = std::sqrt((v[0] * v[0] + v[1] * v[1]) + v[2] * v[2]); v[0] /= l; v[1] /= l; v[2] /= l; };
    auto cross = [](const float a[3], const float b[3], float c[3]) {
        c[0] = a[1] * b[2] - a[2] * b[1]; c[1] = a[2] * b[0] - a[0] * b[2]; c[2] = a[0] * b[1] - a[1] * b[0]; };
    auto dot = [](const float a[3], const float b[3]) { return (a[0] * b[0] + a[1] * b[1]) + a[2] * b[2]; };
    float f[3] = {center[0] - eye[0], center[1] - eye[1], center[2] - eye[2]}; norm(f);
    float s[3]; cross(f, up, s); norm(s);
    float u[3]; cross(s, f, u);
    o[0] = s[0]; o[1] = u[0]; o[2] = -f[0]; o[3] = 0;
    o[4] = s[1]; o[5] = u[1]; o[6] = -f[1]; o[7] = 0;
    o[8] = s[2]; o[9] = u[2]; o[10] = -f[2]; o[11] = 0;
    o[12] = -dot(s, eye); o[13] = -dot(u, eye); o[14] = dot(f, eye); o[15] = 1;
}

/* src/trackball_camera.mm:131-134,150-163 */
void orb_perspective(float fov_deg, int vw, int vh, float zn, float zf, float o[16]) {
    const float aspect = (float)vw / (float)vh;
    const float fovRadians = (float)(fov_deg * M_PI / 180.0f);  /* double expression narrowed by the float parameter */
    const float yScale = 1.0f / std::tan(fovRadians * 0.5f);          /* float overload, as <cmath> resolves it */
    const float xScale = yScale / aspect;
    const float zRange = zf - zn;
    const float zScale = -(zf + zn) / zRange;
    const float wzScale = -2.0f * zf * zn / zRange;
    std::memset(o, 0, 64);
    o[0] = xScale; o[5] = yScale; o[10] = zScale; o[11] = -1.0f; o[14] = wzScale;
}

/* simd_mul(a, b): column j of out = sum_k a.col[k] * b.col[j][k], k = 0..3 */
void orb_mat4_mul(const float a[16], const float b[16], float o[16]) {
    float t[16];
    for (int j = 0; j < 4; ++j) for (int r = 0; r < 4; ++r)
        t[j * 4 + r] = ((a[0 * 4 + r] * b[j * 4 + 0] + a[1 * 4 + r] * b[j * 4 + 1]) + a[2 * 4 + r] * b[j * 4 + 2]) + a[3 * 4 + r] * b[j * 4 + 3];
    std::memcpy(o, t, 64);
}

float orb_exp_neg_half(float q) { return exp_neg_half(q); }
uint16_t orb_float_to_half(float f) { return f2h(f); }
float orb_half_to_float(uint16_t h) { return h2f(h); }

void orb_sort_pairs(uint64_t* keys, uint32_t* vals, uint64_t n) {
    std::vector<uint64_t> k(keys, keys + n); std::vector<uint32_t> v(vals, vals + n);
    lsd_sort(k, v);
    std::memcpy(keys, k.data(), n * 8); std::memcpy(vals, v.data(), n * 4);
}

} // extern "C"
