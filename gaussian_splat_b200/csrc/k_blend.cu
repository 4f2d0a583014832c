// K4 tile ranges and K5 per-16x16-tile blend.
// K5 folds the reference's fragment stage and per-pixel sort+composite into one kernel:
//   TILE    fragment_main + tile_sort_composite   shaders/gaussian_splat_tile.metal:178-267
//   LAYER50 fragment_accumulate + compute_sort_composite
//                                                 shaders/gaussian_splat_50layer.metal:149-223
// Per-pixel fragment lists never touch HBM (the reference keeps 504 B/pixel there and
// clears it every frame, src/instanced_splat_renderer.mm:256-262,475-477).
#include "splat_math.cuh"

// ------------------------------------------------------------------ K4
// bounds[t] = lower_bound(keys, t << 32) for t = 0..T; tile t owns [bounds[t], bounds[t+1]).
// One warp per boundary, 32 probes per round (5 dependent L2 round trips for D < 32M instead of
// the 21+ of a scalar binary search; the kernel is pure latency).
__global__ void __launch_bounds__(256) tile_bounds_kernel(const unsigned long long* __restrict__ keys,
                                                          const FrameCtl* __restrict__ ctl, unsigned int n_bounds,
                                                          unsigned int* __restrict__ bounds)
{
    const unsigned int t = blockIdx.x * 8u + (threadIdx.x >> 5), lane = threadIdx.x & 31u;
    if (t >= n_bounds) return;
    const unsigned int n = ctl->overflow ? 0u : ctl->n_sort;
    const unsigned long long x = (unsigned long long)t << 32;
    unsigned int lo = 0, hi = n;
    while (lo < hi) {
        const unsigned int span = hi - lo;
        const unsigned int pos = lo + (unsigned int)(((unsigned long long)(lane + 1) * span) / 33u);
        const unsigned int c = __popc(__ballot_sync(0xFFFFFFFFu, __ldg(keys + pos) < x));
        const unsigned int new_lo = c == 0 ? lo : lo + (unsigned int)(((unsigned long long)c * span) / 33u) + 1u;
        const unsigned int new_hi = c == 32 ? hi : lo + (unsigned int)(((unsigned long long)(c + 1) * span) / 33u);
        lo = new_lo; hi = new_hi;
    }
    if (lane == 0) bounds[t] = lo;
}

void launch_tile_bounds(const unsigned long long* keys_sorted, const FrameCtl* ctl, unsigned int n_tiles,
                        unsigned int* bounds, cudaStream_t s)
{
    tile_bounds_kernel<<<(n_tiles + 1 + 7) / 8, 256, 0, s>>>(keys_sorted, ctl, n_tiles + 1, bounds);
}

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

template <int MODE>
__global__ void __launch_bounds__(BLEND_THREADS) blend_kernel(
    const __grid_constant__ FrameUniforms U, const splat_record* __restrict__ rec,
    const unsigned int* __restrict__ vals, const unsigned int* __restrict__ bounds, uint2* __restrict__ fb)
{
    __shared__ float4 s_ab[2 * BLEND_BATCH];  // [0..B): cx cy ux uy   [B..2B): vx vy opacity rg
    __shared__ unsigned int s_cm[2 * BLEND_BATCH];  // [0..B): bd   [B..2B): bit b = may touch the 8x4 block of warp b
    float4* const s_a = s_ab; float4* const s_b = s_ab + BLEND_BATCH;
    unsigned int* const s_c = s_cm; unsigned int* const s_m = s_cm + BLEND_BATCH;
    extern __shared__ unsigned int s_list[]; // LAYER50: [K][256] key|vis, [K][256] rg, [K][256] b (as u32)

    const int tx = blockIdx.x, ty = U.row_begin + blockIdx.y;
    const unsigned int tile = (unsigned int)(ty * U.tiles_x + tx);
    int lx, ly; pixel_of_thread(lx, ly);
    const int pi = tx * 16 + lx, pj = ty * 16 + ly;
    const bool inside = pi < U.w && pj < U.h;
    const float pxc = (float)pi + 0.5f, pyc = (float)pj + 0.5f;
    const uint2 range = make_uint2(__ldg(bounds + tile), __ldg(bounds + tile + 1));
    const int K = U.K;
    const unsigned int tid = threadIdx.x;
    const unsigned int a_ab = smem_addr(s_ab), a_cm = smem_addr(s_cm);

    // TILE: (aR, aG) and (aB, accumAlpha) as two half2 accumulators; every half op below is an
    // _rn intrinsic (rounded separately, never contracted: DESIGN.md D7).
    __half2 accRG = __halves2half2(__ushort_as_half(0), __ushort_as_half(0));
    __half2 accBA = accRG;
    const __half h_one = __ushort_as_half(0x3C00);
    // LAYER50 accumulators
    __half aR = __ushort_as_half(0), aG = aR, aB = aR, acc = h_one;
    // cnt counts kept fragments; cnt >= K doubles as "this pixel is finished" (cap reached,
    // alpha saturated, or pixel outside the viewport), so the hot loop carries one counter only.
    int cnt = inside ? 0 : K;
#define done (cnt >= K)

    const unsigned int lane = tid & 31, warp = tid >> 5;
    const float tox = (float)(tx * 16), toy = (float)(ty * 16);
    for (unsigned int b0 = range.x; b0 < range.y; b0 += BLEND_BATCH) {
        const unsigned int nb = min((unsigned int)BLEND_BATCH, range.y - b0);
        unsigned int wmask = 0;
        if (tid < nb) {
            const unsigned int si = __ldg(vals + b0 + tid);
            const float4* r = reinterpret_cast<const float4*>(rec + si);
            const float4 A = __ldg(r), B = __ldg(r + 1);
            s_a[tid] = A;
            s_b[tid] = B;
            s_c[tid] = __float_as_uint(__ldg(reinterpret_cast<const float*>(r + 2)));
            // Conservative pixel AABB of {q <= 9.2103404}: d = Winv (u,v), Winv = [vy -uy; -vx ux] / det.
            // Only used to skip 8x4 pixel blocks the splat cannot touch; never changes a result.
            const float det = A.z * B.y - A.w * B.x;
            const float idet = 1.0f / det;
            const float hx = 3.0351f * sqrtf(B.y * B.y + A.w * A.w) * fabsf(idet) + 0.01f;
            const float hy = 3.0351f * sqrtf(B.x * B.x + A.z * A.z) * fabsf(idet) + 0.01f;
            if (!(hx <= 1e6f && hy <= 1e6f)) wmask = 0xFFu;
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
                unsigned int bits = __ballot_sync(0xFFFFFFFFu, (s_m[c0 + lane] >> warp) & 1u);
                while (bits) {
                    const unsigned int e = c0 + __ffs(bits) - 1;
                    bits &= bits - 1;
                    const float4 A = lds128<0>(a_ab + e * 16u);
                    const float4 B = lds128<BLEND_BATCH * 16>(a_ab + e * 16u);
                    const float dx = pxc - A.x, dy = pyc - A.y;
                    const float u = fmaf(dx, A.z, dy * A.w);
                    const float v = fmaf(dx, B.x, dy * B.y);
                    const float q = fmaf(u, u, v * v);
                    if (done || !(q <= 9.2103404f) || !(fabsf(u) <= 3.0f) || !(fabsf(v) <= 3.0f)) continue;
                    const float alpha = exp_neg_half(q) * B.z;
                    const unsigned int rg = __float_as_uint(B.w), bd = lds32<0>(a_cm + e * 4u);
                    if (MODE == SPLAT_MODE_TILE) {
                        // gaussian_splat_tile.metal:252-264
                        const __half ah = __float2half_rn(alpha);
                        const __half src = __hmul_rn(ah, __hsub_rn(h_one, __high2half(accBA)));
                        const __half2 src2 = __half2half2(src);
                        const unsigned int b1 = (bd & 0xFFFFu) | 0x3C000000u;      // (b, 1.0h): 1.0h * src == src
                        accRG = __hadd2_rn(accRG, __hmul2_rn(*reinterpret_cast<const __half2*>(&rg), src2));
                        accBA = __hadd2_rn(accBA, __hmul2_rn(*reinterpret_cast<const __half2*>(&b1), src2));
                        cnt = __hge(__high2half(accBA), __ushort_as_half(0x3BEC)) ? K : cnt + 1;
                    } else {
                        // gaussian_splat_50layer.metal:170-176: first K arrivals (instance order) kept
                        const unsigned int h = bd >> 16;
                        const unsigned int dk = U.order == SPLAT_ORDER_REFERENCE ? (0xFFFFu - h) : h;
                        const unsigned int vis = hbits(__float2half_rn(1.0f - alpha));
                        s_list[(0 * K + cnt) * BLEND_THREADS + tid] = (dk << 16) | vis;
                        s_list[(1 * K + cnt) * BLEND_THREADS + tid] = rg;
                        s_list[(2 * K + cnt) * BLEND_THREADS + tid] = bd & 0xFFFFu;
                        ++cnt;
                    }
                }
                if (__all_sync(0xFFFFFFFFu, done)) break;
            }
        }
        if (__syncthreads_and(done)) break;
    }

#undef done
    if (!inside) return;
    uint2 out;
    if (MODE == SPLAT_MODE_TILE) {
        out.x = *reinterpret_cast<const unsigned int*>(&accRG);
        out.y = *reinterpret_cast<const unsigned int*>(&accBA);
    } else if (cnt == 0) {
        out = make_uint2(0, 0);
    } else {
        // gaussian_splat_50layer.metal:198-220. The stable descending insertion sort equals an
        // ascending order on (depth key, arrival slot); walk it by repeated selection so an
        // early exit (accumVis < 0.01h) skips the rest of the sort.
        unsigned int last = 0;      // (dk << 8 | slot) + 1 of the previous pick
        for (int step = 0; step < cnt; ++step) {
            unsigned int best = 0xFFFFFFFFu;
            for (int s = 0; s < cnt; ++s) {
                const unsigned int kk = ((s_list[(0 * K + s) * BLEND_THREADS + tid] >> 16) << 8 | (unsigned int)s) + 1u;
                if (kk > last && kk < best) best = kk;
            }
            last = best;
            const int s = (int)((best - 1u) & 0xFFu);
            const unsigned int w0 = s_list[(0 * K + s) * BLEND_THREADS + tid];
            const unsigned int rg = s_list[(1 * K + s) * BLEND_THREADS + tid];
            const unsigned int bb = s_list[(2 * K + s) * BLEND_THREADS + tid];
            aR = __hadd_rn(aR, __hmul_rn(__ushort_as_half((unsigned short)(rg & 0xFFFFu)), acc));
            aG = __hadd_rn(aG, __hmul_rn(__ushort_as_half((unsigned short)(rg >> 16)), acc));
            aB = __hadd_rn(aB, __hmul_rn(__ushort_as_half((unsigned short)bb), acc));
            acc = __hmul_rn(acc, __ushort_as_half((unsigned short)(w0 & 0xFFFFu)));
            if (__hlt(acc, __ushort_as_half(0x211F))) break;
        }
        out.x = (unsigned int)hbits(aR) | ((unsigned int)hbits(aG) << 16);
        out.y = (unsigned int)hbits(aB) | ((unsigned int)hbits(__hsub_rn(h_one, acc)) << 16);
    }
    fb[(size_t)pj * U.w + pi] = out;
}

int blend_layer_smem_bytes(int K) { return 3 * K * BLEND_THREADS * 4; }

void launch_blend(const FrameUniforms& U, const splat_record* rec, const unsigned int* vals_sorted,
                  const unsigned int* bounds, void* fb_half4, cudaStream_t s)
{
    const int rows = U.row_end - U.row_begin;
    if (rows <= 0 || U.tiles_x <= 0) return;
    const dim3 grid(U.tiles_x, rows);
    if (U.mode == SPLAT_MODE_LAYER50) {
        const int smem = blend_layer_smem_bytes(U.K);
        cudaFuncSetAttribute(blend_kernel<SPLAT_MODE_LAYER50>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        blend_kernel<SPLAT_MODE_LAYER50><<<grid, BLEND_THREADS, smem, s>>>(U, rec, vals_sorted, bounds,
                                                                          reinterpret_cast<uint2*>(fb_half4));
    } else {
        blend_kernel<SPLAT_MODE_TILE><<<grid, BLEND_THREADS, 0, s>>>(U, rec, vals_sorted, bounds,
                                                                     reinterpret_cast<uint2*>(fb_half4));
    }
}
