// K2: exclusive scan of tiles_touched (tile prefixes come from K1's last CTA, so there is no ticket and
// no lookback chain here) fused with (tile | depth) key / value duplication and the radix-sort digit histograms.
// Replaces the hardware tiler's binning of the instanced quads
// (src/instanced_splat_renderer.mm:496-499 draw call).
#include "splat_internal.cuh"

__device__ __forceinline__ unsigned int sat_add(unsigned int a, unsigned int b) {
    const unsigned long long s = (unsigned long long)a + b;
    return s > 0xFFFFFFFFull ? 0xFFFFFFFFu : (unsigned int)s;
}

// Adds 1 to hist[digit] for every active lane, one shared-memory atomic per distinct digit.
__device__ __forceinline__ void warp_hist_inc(unsigned int* hist, unsigned int digit) {
    const unsigned int act = __activemask();
    const unsigned int peers = __match_any_sync(act, digit);
    if ((peers & ((1u << (threadIdx.x & 31)) - 1u)) == 0) atomicAdd(hist + digit, __popc(peers));
}

// Persistent grid: a CTA walks duplication tiles (1024 splats) with a fixed stride and keeps ONE digit
// histogram for all of them, flushed once at the end. With one CTA per tile, a 6M-splat band frame issued
// 5860 x up to 1024 same-address global atomics (a band has 2-3 distinct high tile digits); now at most
// grid x 1024. Tiles that produced no pair (K1 left their pair count) only write their offsets.
__global__ void __launch_bounds__(SCAN_THREADS) scan_duplicate_kernel(
    const __grid_constant__ FrameUniforms U, const splat_record* __restrict__ rec,
    const unsigned int* __restrict__ touched, unsigned int* __restrict__ offsets,
    unsigned long long* __restrict__ keys, unsigned int* __restrict__ vals,
    unsigned long long* __restrict__ keys_depth, unsigned int* __restrict__ vals_depth, unsigned long long capacity,
    FrameCtl* ctl, const unsigned int* __restrict__ tile_sums, const unsigned int* __restrict__ scan_prefix,
    const __grid_constant__ SortPlan plan, unsigned int n_tiles, unsigned int n_k1_tiles)
{
    __shared__ unsigned int s_scan[SCAN_TILE_ITEMS];
    __shared__ unsigned int s_rlo[SCAN_TILE_ITEMS], s_rhi[SCAN_TILE_ITEMS], s_df[SCAN_TILE_ITEMS];
    __shared__ unsigned int s_hist[FRAME_MAX_PASSES][256];
    __shared__ unsigned int s_warp[SCAN_THREADS / 32];

    const unsigned int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int p = 0; p < FRAME_MAX_PASSES; ++p) s_hist[p][tid] = 0;
    int tile_pass[2], n_tile_pass = 0;
    for (int p = 0; p < plan.n_pass; ++p) if (plan.shift[p] >= 32 && n_tile_pass < 2) tile_pass[n_tile_pass++] = p;
    const bool layer = U.mode == SPLAT_MODE_LAYER50;        // second, depth-ordered list
    const bool arrival = U.mode != SPLAT_MODE_TILE;         // LAYER50 / MLAB: the list the pixels walk is in instance order
    __syncthreads();

    for (unsigned int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const unsigned int base = tile * SCAN_TILE_ITEMS;
        const unsigned int prefix = __ldg(scan_prefix + tile);      // saturated at 2^32 - 1, like the offsets
        const unsigned int btotal = __ldg(tile_sums + 2 * tile) + (2 * tile + 1 < n_k1_tiles ? __ldg(tile_sums + 2 * tile + 1) : 0u);
        if (btotal == 0) {      // nothing visible in this tile of splats (common with tile-row bands)
            *reinterpret_cast<uint4*>(offsets + base + tid * SCAN_ITEMS) = make_uint4(prefix, prefix, prefix, prefix);
            continue;
        }
        // ---- load 4 consecutive counts per thread (buffers are padded to the tile size)
        const uint4 c4 = *reinterpret_cast<const uint4*>(touched + base + tid * SCAN_ITEMS);
        unsigned int cnt[SCAN_ITEMS] = {c4.x, c4.y, c4.z, c4.w};
        unsigned int tsum = 0;
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; ++k) tsum += cnt[k];
        unsigned int incl = tsum;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { unsigned int v = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= d) incl += v; }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        unsigned int wbase = 0;
#pragma unroll
        for (int k = 0; k < SCAN_THREADS / 32; ++k) { unsigned int v = s_warp[k]; if (k < (int)warp) wbase += v; }
        unsigned int excl = wbase + incl - tsum;   // thread-exclusive, block-local

        // ---- stage per-splat rect / depth field; per-splat (depth-digit) histograms; offsets
        uint4 o;
        unsigned int* ov = &o.x;
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; ++k) {
            const unsigned int li = tid * SCAN_ITEMS + k;
            s_scan[li] = excl;
            ov[k] = sat_add(prefix, excl);
            excl += cnt[k];
            if (cnt[k]) {
                const float4 tail = __ldg(reinterpret_cast<const float4*>(rec + base + li) + 2);  // bd, z, rect_lo, rect_hi
                const unsigned int h = __float_as_uint(tail.x) >> 16;
                const unsigned int df = U.order == SPLAT_ORDER_REFERENCE ? (0xFFFFu - h) : h;   // half-depth field (D3)
                s_df[li] = df;
                s_rlo[li] = __float_as_uint(tail.z);
                s_rhi[li] = __float_as_uint(tail.w);
                for (int p = 0; p < plan.n_pass; ++p)
                    if (plan.shift[p] < 32) atomicAdd(&s_hist[plan.slot[p]][(df >> plan.shift[p]) & plan.mask[p]], cnt[k]);
            }
        }
        *reinterpret_cast<uint4*>(offsets + base + tid * SCAN_ITEMS) = o;
        __syncthreads();

        // ---- cooperative expansion: output slot j -> (splat, k-th tile of its rectangle)
        for (unsigned int j = tid; j < btotal; j += SCAN_THREADS) {
            unsigned int lo = 0, hi = SCAN_TILE_ITEMS;
            while (lo < hi) { const unsigned int mid = (lo + hi) >> 1; if (s_scan[mid] <= j) lo = mid + 1; else hi = mid; }
            const unsigned int sidx = lo - 1;
            const unsigned int k = j - s_scan[sidx];
            const unsigned int rlo = s_rlo[sidx], rhi = s_rhi[sidx];
            const unsigned int tx0 = rlo & 0xFFFFu, ty0 = rlo >> 16, tx1 = rhi & 0xFFFFu;
            const unsigned int wdt = tx1 - tx0 + 1;
            const unsigned int ry = k / wdt, rx = k - ry * wdt;
            const unsigned int t = (ty0 + ry) * (unsigned int)U.tiles_x + (tx0 + rx);
            for (int q = 0; q < n_tile_pass; ++q)
                warp_hist_inc(s_hist[plan.slot[tile_pass[q]]], (t >> (plan.shift[tile_pass[q]] - 32)) & plan.mask[tile_pass[q]]);
            const unsigned long long g = (unsigned long long)prefix + j;
            if (g < capacity) {
                // TILE: tile | depth. LAYER50 / MLAB: tile | instance index (arrival order made explicit); LAYER50
                // adds a second copy in tile | depth order for the composite.
                keys[g] = ((unsigned long long)t << 32) | (arrival ? base + sidx : s_df[sidx]);
                vals[g] = base + sidx;
                if (layer) {
                    keys_depth[g] = ((unsigned long long)t << 32) | s_df[sidx];
                    vals_depth[g] = base + sidx;
                }
            }
        }
        __syncthreads();            // s_scan / s_rlo / s_rhi / s_df / s_warp are reused by the next tile
    }
    __syncthreads();
    for (int p = 0; p < FRAME_MAX_PASSES; ++p) {
        const unsigned int v = s_hist[p][tid];
        if (v) atomicAdd(&ctl->hist[p][tid], v);
    }
}

void launch_scan_duplicate(const FrameUniforms& U, const splat_record* rec, const unsigned int* touched,
                           unsigned int* offsets, unsigned long long* keys, unsigned int* vals,
                           unsigned long long* keys_depth, unsigned int* vals_depth,
                           unsigned long long capacity, FrameCtl* ctl, const unsigned int* tile_sums,
                           const unsigned int* scan_prefix, SortPlan plan, int sm_count, cudaStream_t s)
{
    if (U.n == 0) return;
    const unsigned int n_tiles = (U.n + SCAN_TILE_ITEMS - 1) / SCAN_TILE_ITEMS;
    const unsigned int n_k1 = (U.n + SCAN_TILE_ITEMS / 2 - 1) / (SCAN_TILE_ITEMS / 2);
    const unsigned int grid = n_tiles < (unsigned int)sm_count * 8u ? n_tiles : (unsigned int)sm_count * 8u;
    scan_duplicate_kernel<<<grid, SCAN_THREADS, 0, s>>>(U, rec, touched, offsets, keys, vals, keys_depth, vals_depth, capacity, ctl,
                                                        tile_sums, scan_prefix, plan, n_tiles, n_k1);
}
