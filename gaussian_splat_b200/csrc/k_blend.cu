// K5: per-16x16-tile blend (the tile lists and their bounds come from k_bin.cu).
// K5 folds the reference's fragment stage and per-pixel sort+composite into one kernel:
//   TILE    fragment_main + tile_sort_composite   shaders/gaussian_splat_tile.metal:178-267
//   LAYER50 fragment_accumulate + compute_sort_composite
//                                                 shaders/gaussian_splat_50layer.metal:149-223
//   MLAB    fragment_main (6-layer K-buffer insert + merge) + resolve_main
//                                                 shaders/gaussian_splat.metal:201-291, 329-361
// Per-pixel fragment lists never touch HBM (the reference keeps 504 B/pixel there and
// clears it every frame, src/instanced_splat_renderer.mm:256-262,475-477).
#include "splat_math.cuh"

// ------------------------------------------------------------------ K5
#define BLEND_THREADS 256
#define BLEND_BATCH 256

__device__ __forceinline__ unsigned short hbits(__half h) { return __half_as_ushort(h); }

// Thread -> pixel mapping: a warp covers an 8x4 pixel block (better splat coherence than 16x2).
__device__ __forceinline__ void pixel_of_thread(int& lx, int& ly) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    lx = (warp & 1) * 8 + (lane & 7);
    ly = (warp >> 1) * 4 + (lane >> 3);
}

// Shared-memory access by 32-bit shared-window address kept in a register: the compiler
// otherwise rematerialises the window base (S2R SR_CgaCtaId + LEA) inside the hot loop.
__device__ __forceinline__ unsigned int smem_addr(const void* p) {
    unsigned int a = (unsigned int)__cvta_generic_to_shared(p), o;
    asm volatile("mov.u32 %0, %1;" : "=r"(o) : "r"(a));     // opaque: not rematerialisable
    return o;
}
template <int OFS>
__device__ __forceinline__ float4 lds128(unsigned int a) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4+%5];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a), "n"(OFS));
    return v;
}
template <int OFS>
__device__ __forceinline__ unsigned int lds32(unsigned int a) {
    unsigned int v;
    asm volatile("ld.shared.u32 %0, [%1+%2];" : "=r"(v) : "r"(a), "n"(OFS));
    return v;
}

// What the walker does with a covering fragment:
//   WALK_TILE     composite in list order                         gaussian_splat_tile.metal:252-264
//   WALK_COUNT    LAYER50 phase 1: count arrivals (instance order), remember the K-th arrival
//                                                                  gaussian_splat_50layer.metal:170-176
//   WALK_UNDER    LAYER50 phase 2: "under" composite in depth order of the arrivals that were kept
//                                                                  gaussian_splat_50layer.metal:209-220
//   WALK_MLAB     insert into / merge with the 6-layer K-buffer in arrival (instance) order
//                                                                  gaussian_splat.metal:221-276
enum { WALK_TILE = 0, WALK_COUNT = 1, WALK_UNDER = 2, WALK_MLAB = 3 };

// MLAB K-buffer of one pixel, in registers: six (premultiplied rgb, visibility) layers and their half depths.
// The reference keeps them in eight RGBA16F attachments cleared to (0,0,0,1)
// (src/instanced_splat_renderer.mm:536-541), so depth slot 3 - depths01.a - starts at 1.0h, the others at 0.
struct KBuffer {
    __half2 rg[6], ba[6];
    __half d[6];
};

struct PixelState {
    __half2 accRG, accBA;     // (aR, aG) and (aB, accumAlpha | accumVis): every op an _rn intrinsic (D7)
    int rem;                  // fragments the pixel can still take; <= 0 doubles as "this pixel is finished"
    unsigned int cutoff;      // LAYER50: instance index of the K-th arrival (fragments after it were dropped)
};

// Walks one tile's pair list [range.x, range.y) in batches of BLEND_BATCH staged through shared
// memory. Each staging thread also derives a conservative pixel AABB of its splat and an 8-bit mask
// of the 8x4-pixel warp blocks it can touch; each warp ballots the masks and visits only its own
// candidates. Returns when every pixel of the tile is finished or the list ends.
// vals[VSTRIDE * e] = instance index of list entry e (VSTRIDE 2: the low words of the 64-bit pairs, 1: LAYER50's arrival list).
// rec / vals are written earlier in the same frame by kernels this one is linked to by programmatic dependent launch:
// they are read through L2 (ld.global.cg), not through the non-coherent path.
template <int WALK, int VSTRIDE, bool COUNT>
__device__ __forceinline__ void walk_list(const splat_record_row* rec, const unsigned int* vals,
                                          uint2 range, int K, float pxc, float pyc, float tox, float toy,
                                          float4* s_ab, unsigned int* s_cm, PixelState& px, KBuffer& kb, unsigned int& n_cand)
{
    float4* const s_a = s_ab; float4* const s_b = s_ab + BLEND_BATCH;
    unsigned int* const s_c = s_cm; unsigned int* const s_m = s_cm + BLEND_BATCH; unsigned int* const s_i = s_cm + 2 * BLEND_BATCH;
    const unsigned int a_ab = smem_addr(s_ab), a_cm = smem_addr(s_cm);
    const unsigned int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const __half h_one = __ushort_as_half(0x3C00);
    float exp_c5;
    asm volatile("mov.f32 %0, 0f3AADDD2A;" : "=f"(exp_c5));      // EXP_NEG_HALF_C5, pinned to a register
#define done (px.rem <= 0)
    for (unsigned int b0 = range.x; b0 < range.y; b0 += BLEND_BATCH) {
        const unsigned int nb = min((unsigned int)BLEND_BATCH, range.y - b0);
        unsigned int wmask = 0;
        if (tid < nb) {
            const unsigned int si = __ldcg(vals + (size_t)(b0 + tid) * VSTRIDE);
            const float4* r = reinterpret_cast<const float4*>(rec + si);
            const float4 A = __ldcg(r), B = __ldcg(r + 1);
            s_a[tid] = A;
            s_b[tid] = B;
            if (WALK != WALK_COUNT) {
                // TILE: (b, 1.0h) so that one half2 multiply gives (b * src, src); UNDER: (b, 0); MLAB: (b, depth) as stored
                const unsigned int bd = __float_as_uint(__ldcg(reinterpret_cast<const float*>(r + 2)));
                s_c[tid] = WALK == WALK_TILE ? ((bd & 0xFFFFu) | 0x3C000000u) : WALK == WALK_UNDER ? (bd & 0xFFFFu) : bd;
            }
            const bool too_near = WALK == WALK_MLAB && __ldcg(reinterpret_cast<const float*>(r + 2) + 1) < 0.001f;   // gaussian_splat.metal:209
            if (WALK != WALK_TILE) s_i[tid] = si;
            // Conservative pixel AABB of {q <= 9.2103404}: d = Winv (u,v), Winv = [vy -uy; -vx ux] / det.
            // Only used to skip 8x4 pixel blocks the splat cannot touch; never changes a result.
            const float det = A.z * B.y - A.w * B.x;
            const float idet = 1.0f / det;
            const float hx = 3.0351f * sqrtf(B.y * B.y + A.w * A.w) * fabsf(idet) + 0.01f;
            const float hy = 3.0351f * sqrtf(B.x * B.x + A.z * A.z) * fabsf(idet) + 0.01f;
            if (too_near) wmask = 0;
            else if (!(hx <= 1e6f && hy <= 1e6f)) wmask = 0xFFu;
            else {
                const float x0 = A.x - hx - tox, x1 = A.x + hx - tox, y0 = A.y - hy - toy, y1 = A.y + hy - toy;
                unsigned int cm = 0, rm = 0;
                if (x1 >= 0.5f && x0 <= 7.5f) cm |= 1u;
                if (x1 >= 8.5f && x0 <= 15.5f) cm |= 2u;
#pragma unroll
                for (int by = 0; by < 4; ++by)
                    if (y1 >= 4.0f * by + 0.5f && y0 <= 4.0f * by + 3.5f) rm |= 1u << by;
#pragma unroll
                for (int by = 0; by < 4; ++by)
                    if (rm & (1u << by)) wmask |= cm << (2 * by);
            }
        }
        s_m[tid] = wmask;
        __syncthreads();
        if (!__all_sync(0xFFFFFFFFu, done)) {
            for (unsigned int c0 = 0; c0 < nb; c0 += 32) {
                // candidates of this warp; lane j looks at entry c0 + 31 - j, so the FIRST entry is the HIGHEST
                // set bit: one bfind per step instead of the brev + bfind of a find-first-set
                unsigned int bits = __ballot_sync(0xFFFFFFFFu, (s_m[c0 + 31u - lane] >> warp) & 1u);
                const unsigned int a_top = a_ab + (c0 + 31u) * 16u, c_top = a_cm + (c0 + 31u) * 4u;
                while (bits) {
                    unsigned int j;
                    if (COUNT) ++n_cand;            // warp-candidates executed (SPLAT_FLAG_COUNT_WORK: the roofline's work count)
                    asm("bfind.u32 %0, %1;" : "=r"(j) : "r"(bits));
                    {   // clear the top bit: keep bits [0, j)
                        unsigned int below;
                        asm("bmsk.clamp.b32 %0, %1, %2;" : "=r"(below) : "r"(0), "r"(j));
                        bits &= below;
                    }
                    const unsigned int a_e = a_top - j * 16u, c_e = c_top - j * 4u;
                    const float4 A = lds128<0>(a_e);
                    const float4 B = lds128<BLEND_BATCH * 16>(a_e);
                    const float dx = pxc - A.x, dy = pyc - A.y;
                    const float u = fmaf(dx, A.z, dy * A.w);
                    const float v = fmaf(dx, B.x, dy * B.y);
                    const float q = fmaf(u, u, v * v);
                    if (done || !(q <= 9.2103404f) || !(fabsf(u) <= 3.0f) || !(fabsf(v) <= 3.0f)) continue;
                    if (WALK == WALK_COUNT) {
                        // arrival number cnt+1; the K-th closes the pixel's list
                        if (--px.rem <= 0) px.cutoff = lds32<2 * BLEND_BATCH * 4>(c_e);
                        continue;
                    }
                    if (WALK == WALK_UNDER && lds32<2 * BLEND_BATCH * 4>(c_e) > px.cutoff) continue;   // arrived after the cap
                    const float alpha = exp_neg_half(q, exp_c5) * B.z;
                    const unsigned int rg = __float_as_uint(B.w), bd = lds32<0>(c_e);
                    if (WALK == WALK_MLAB) {
                        // newLayer = half4(half3(color) * half(alpha), 1h - half(alpha)); newDepth = half(depth)
                        const __half ah = __float2half_rn(alpha);
                        __half2 nrg = __hmul2_rn(*reinterpret_cast<const __half2*>(&rg), __half2half2(ah));
                        __half2 nba = __halves2half2(__hmul_rn(__ushort_as_half((unsigned short)(bd & 0xFFFFu)), ah), __hsub_rn(h_one, ah));
                        __half nd = __ushort_as_half((unsigned short)(bd >> 16));
#pragma unroll
                        for (int k = 0; k < 6; ++k) {            // insertion: every slot the fragment passes shifts down
                            const bool ins = __hge(nd, kb.d[k]);
                            const __half2 trg = kb.rg[k], tba = kb.ba[k]; const __half td = kb.d[k];
                            kb.rg[k] = ins ? nrg : trg; kb.ba[k] = ins ? nba : tba; kb.d[k] = ins ? nd : td;
                            nrg = ins ? trg : nrg; nba = ins ? tba : nba; nd = ins ? td : nd;
                        }
                        // what fell out merges with the last layer: front UNDER back
                        const bool closer = __hge(nd, kb.d[5]);
                        const __half2 frg = closer ? nrg : kb.rg[5], fba = closer ? nba : kb.ba[5];
                        const __half2 brg = closer ? kb.rg[5] : nrg, bba = closer ? kb.ba[5] : nba;
                        const __half2 ba2 = __half2half2(__high2half(bba));
                        const __half2 t = __hmul2_rn(fba, ba2);                       // (front.b * back.a, front.a * back.a)
                        kb.rg[5] = __hadd2_rn(brg, __hmul2_rn(frg, ba2));
                        kb.ba[5] = __halves2half2(__hadd_rn(__low2half(bba), __low2half(t)), __high2half(t));
                        kb.d[5] = closer ? nd : kb.d[5];
                        continue;
                    }
                    if (WALK == WALK_TILE) {
                        const __half ah = __float2half_rn(alpha);
                        const __half src = __hmul_rn(ah, __hsub_rn(h_one, __high2half(px.accBA)));
                        const __half2 src2 = __half2half2(src);
                        const unsigned int b1 = bd;                                // staged as (b, 1.0h): 1.0h * src == src
                        px.accRG = __hadd2_rn(px.accRG, __hmul2_rn(*reinterpret_cast<const __half2*>(&rg), src2));
                        px.accBA = __hadd2_rn(px.accBA, __hmul2_rn(*reinterpret_cast<const __half2*>(&b1), src2));
                        px.rem = __hge(__high2half(px.accBA), __ushort_as_half(0x3BEC)) ? 0 : px.rem - 1;
                    } else {
                        // accumColor += color * accumVis; accumVis *= visibility; stop below 0.01h.
                        // visibility = half(1 - alpha), float subtract then round (50layer.metal:174).
                        const __half vis = __high2half(px.accBA);
                        const __half2 vis2 = __half2half2(vis);
                        const unsigned int b0h = bd;                              // staged as (b, 0): the alpha lane is replaced below
                        px.accRG = __hadd2_rn(px.accRG, __hmul2_rn(*reinterpret_cast<const __half2*>(&rg), vis2));
                        const __half2 bsum = __hadd2_rn(px.accBA, __hmul2_rn(*reinterpret_cast<const __half2*>(&b0h), vis2));
                        const __half nvis = __hmul_rn(vis, __float2half_rn(1.0f - alpha));
                        px.accBA = __halves2half2(__low2half(bsum), nvis);
                        if (__hlt(nvis, __ushort_as_half(0x211F))) px.rem = 0;
                    }
                }
                if (__all_sync(0xFFFFFFFFu, done)) break;
            }
        }
        if (__syncthreads_and(done)) break;
    }
#undef done
}

template <int MODE, bool COUNT>
__global__ void __launch_bounds__(BLEND_THREADS) blend_kernel(
    const __grid_constant__ FrameUniforms U, const splat_record_row* rec,
    const unsigned int* pairs_lo, const unsigned int* arrival,
    const unsigned int* bounds, const unsigned int* order, uint2* __restrict__ fb, FrameCtl* ctl)
{
    __shared__ float4 s_ab[2 * BLEND_BATCH];        // [0..B): cx cy ux uy   [B..2B): vx vy opacity rg
    __shared__ unsigned int s_cm[3 * BLEND_BATCH];  // [0..B): bd   [B..2B): 8x4-block mask   [2B..3B): instance index
    pdl_prologue();

    const unsigned int tile = __ldcg(order + blockIdx.x);          // the band's tiles, longest lists first (tile_offsets_kernel)
    const int ty = (int)(tile / (unsigned int)U.tiles_x), tx = (int)(tile - (unsigned int)ty * (unsigned int)U.tiles_x);
    int lx, ly; pixel_of_thread(lx, ly);
    const int pi = tx * 16 + lx, pj = ty * 16 + ly;
    const bool inside = pi < U.w && pj < U.h;
    const float pxc = (float)pi + 0.5f, pyc = (float)pj + 0.5f;
    const float tox = (float)(tx * 16), toy = (float)(ty * 16);
    const uint2 range = make_uint2(__ldcg(bounds + tile), __ldcg(bounds + tile + 1));
    const int K = U.K;

    PixelState px;
    KBuffer kb;
    unsigned int n_cand = 0;
    px.accRG = __halves2half2(__ushort_as_half(0), __ushort_as_half(0));
    px.cutoff = 0xFFFFFFFFu;
    if (MODE == SPLAT_MODE_TILE) {
        px.accBA = px.accRG;
        px.rem = inside ? K : 0;
        walk_list<WALK_TILE, 2, COUNT>(rec, pairs_lo, range, K, pxc, pyc, tox, toy, s_ab, s_cm, px, kb, n_cand);
    } else if (MODE == SPLAT_MODE_MLAB) {
        const __half h0 = __ushort_as_half(0), h1 = __ushort_as_half(0x3C00);
#pragma unroll
        for (int k = 0; k < 6; ++k) { kb.rg[k] = __halves2half2(h0, h0); kb.ba[k] = __halves2half2(h0, h1); kb.d[k] = k == 3 ? h1 : h0; }
        px.accBA = px.accRG;
        px.rem = inside ? K : 0;                       // K = INT_MAX: a pixel never closes
        walk_list<WALK_MLAB, 2, COUNT>(rec, pairs_lo, range, K, pxc, pyc, tox, toy, s_ab, s_cm, px, kb, n_cand);
        // resolve_main: composite the six layers front to back
        __half2 crg = __halves2half2(h0, h0); __half cb = h0, at = h1;
#pragma unroll
        for (int k = 0; k < 6; ++k) {
            crg = __hadd2_rn(crg, __hmul2_rn(kb.rg[k], __half2half2(at)));
            cb = __hadd_rn(cb, __hmul_rn(__low2half(kb.ba[k]), at));
            at = __hmul_rn(at, __high2half(kb.ba[k]));
        }
        px.accRG = crg;
        px.accBA = __halves2half2(cb, __hsub_rn(h1, at));
    } else {
        // Phase 1 (arrival = instance order): only a counter per pixel; a tile with at most K pairs cannot
        // overflow any list. Phase 2 (depth order, ties in arrival order because the sort is stable):
        // composite the kept arrivals front to back with the shader's early exit. The per-pixel lists of
        // the reference (504 B/pixel in device memory) are never materialised.
        if (range.y - range.x > (unsigned int)K) {
            px.rem = inside ? K : 0;
            walk_list<WALK_COUNT, 1, COUNT>(rec, arrival, range, K, pxc, pyc, tox, toy, s_ab, s_cm, px, kb, n_cand);
            __syncthreads();
        }
        px.accBA = __halves2half2(__ushort_as_half(0), __ushort_as_half(0x3C00));   // accumVis = 1
        px.rem = inside ? K : 0;
        walk_list<WALK_UNDER, 2, COUNT>(rec, pairs_lo, range, K, pxc, pyc, tox, toy, s_ab, s_cm, px, kb, n_cand);
    }
    if (COUNT && (threadIdx.x & 31) == 0 && n_cand) atomicAdd(&ctl->blend_warp_candidates, (unsigned long long)n_cand);
    if (!inside) return;
    uint2 out;
    out.x = *reinterpret_cast<const unsigned int*>(&px.accRG);
    if (MODE != SPLAT_MODE_LAYER50) {
        out.y = *reinterpret_cast<const unsigned int*>(&px.accBA);
    } else {
        // half4(accumColor, 1h - accumVis); an empty pixel gives (0,0,0,0) either way (50layer.metal:192-195,222)
        const __half a = __hsub_rn(__ushort_as_half(0x3C00), __high2half(px.accBA));
        out.y = (unsigned int)hbits(__low2half(px.accBA)) | ((unsigned int)hbits(a) << 16);
    }
    fb[(size_t)pj * U.w + pi] = out;
}

template <bool COUNT>
static void launch_blend_t(const FrameUniforms& U, const splat_record_row* rec, const unsigned int* lo, const unsigned int* arrival,
                           const unsigned int* bounds, const unsigned int* order, void* fb_half4, FrameCtl* ctl, cudaStream_t s)
{
    const dim3 grid((unsigned int)(U.tiles_x * (U.row_end - U.row_begin)));
    if (U.mode == SPLAT_MODE_MLAB)
        launch_pdl(blend_kernel<SPLAT_MODE_MLAB, COUNT>, grid, dim3(BLEND_THREADS), 0, s, U, rec, lo, arrival, bounds, order, reinterpret_cast<uint2*>(fb_half4), ctl);
    else if (U.mode == SPLAT_MODE_LAYER50)
        launch_pdl(blend_kernel<SPLAT_MODE_LAYER50, COUNT>, grid, dim3(BLEND_THREADS), 0, s, U, rec, lo, arrival, bounds, order, reinterpret_cast<uint2*>(fb_half4), ctl);
    else
        launch_pdl(blend_kernel<SPLAT_MODE_TILE, COUNT>, grid, dim3(BLEND_THREADS), 0, s, U, rec, lo, arrival, bounds, order, reinterpret_cast<uint2*>(fb_half4), ctl);
}

void launch_blend(const FrameUniforms& U, const splat_record_row* rec, const unsigned long long* pairs,
                  const unsigned int* arrival, const unsigned int* bounds, const unsigned int* order, void* fb_half4, FrameCtl* ctl,
                  bool count_work, cudaStream_t s)
{
    const int rows = U.row_end - U.row_begin;
    if (rows <= 0 || U.tiles_x <= 0) return;
    const unsigned int* lo = reinterpret_cast<const unsigned int*>(pairs);      // little endian: the instance index is word 0 of a pair
    if (count_work) launch_blend_t<true>(U, rec, lo, arrival, bounds, order, fb_half4, ctl, s);
    else launch_blend_t<false>(U, rec, lo, arrival, bounds, order, fb_half4, ctl, s);
}
