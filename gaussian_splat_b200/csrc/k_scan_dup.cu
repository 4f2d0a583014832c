// K2: single-pass decoupled-lookback exclusive scan of tiles_touched fused with
// (tile | depth) key / value duplication and the radix-sort digit histograms.
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

__global__ void __launch_bounds__(SCAN_THREADS) scan_duplicate_kernel(
    const __grid_constant__ FrameUniforms U, const splat_record* __restrict__ rec,
    const unsigned int* __restrict__ touched, unsigned int* __restrict__ offsets,
    unsigned long long* __restrict__ keys, unsigned int* __restrict__ vals,
    unsigned long long* __restrict__ keys_depth, unsigned int* __restrict__ vals_depth, unsigned long long capacity,
    FrameCtl* ctl, unsigned long long* status, unsigned int epoch, const __grid_constant__ SortPlan plan,
    unsigned int n_tiles)
{
    __shared__ unsigned int s_scan[SCAN_TILE_ITEMS];
    __shared__ unsigned int s_rlo[SCAN_TILE_ITEMS], s_rhi[SCAN_TILE_ITEMS], s_df[SCAN_TILE_ITEMS];
    __shared__ unsigned int s_hist[FRAME_MAX_PASSES][256];
    __shared__ unsigned int s_warp[SCAN_THREADS / 32];
    __shared__ unsigned int s_tile, s_prefix;

    const unsigned int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(&ctl->scan_ticket, 1u);
    for (int p = 0; p < FRAME_MAX_PASSES; ++p) s_hist[p][tid] = 0;
    __syncthreads();
    const unsigned int tile = s_tile;
    const unsigned int base = tile * SCAN_TILE_ITEMS;

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
    unsigned int wbase = 0, btotal = 0;
#pragma unroll
    for (int k = 0; k < SCAN_THREADS / 32; ++k) { unsigned int v = s_warp[k]; if (k < (int)warp) wbase += v; btotal += v; }
    unsigned int excl = wbase + incl - tsum;   // thread-exclusive, block-local

    // ---- stage per-splat rect / depth field; per-splat (depth-digit) histograms
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) {
        const unsigned int li = tid * SCAN_ITEMS + k;
        s_scan[li] = excl;
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

    // ---- decoupled lookback (warp 0)
    if (warp == 0) {
        if (lane == 0) lb_store(status + tile, lb_pack(epoch, tile == 0 ? LB_FLAG_INC : LB_FLAG_AGG, btotal));
        unsigned int prefix = 0;
        int look = (int)tile - 1;
        while (look >= 0) {
            const int idx = look - (int)lane;
            unsigned long long wd = lb_pack(epoch, LB_FLAG_INC, 0);
            if (idx >= 0) {
                do { wd = lb_load(status + idx); } while ((unsigned int)(wd >> 34) != epoch || ((wd >> 32) & 3ull) == 0);
            }
            const unsigned int inc_mask = __ballot_sync(0xFFFFFFFFu, ((wd >> 32) & 3ull) == LB_FLAG_INC);
            const int first = __ffs(inc_mask) - 1;      // nearest predecessor holding an inclusive prefix
            unsigned int v = ((int)lane <= first || first < 0) ? (unsigned int)wd : 0u;
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) v = sat_add(v, __shfl_xor_sync(0xFFFFFFFFu, v, d));
            prefix = sat_add(prefix, v);
            if (first >= 0) break;
            look -= 32;
        }
        if (lane == 0) {
            const unsigned int inclusive = sat_add(prefix, btotal);
            if (tile != 0) lb_store(status + tile, lb_pack(epoch, LB_FLAG_INC, inclusive));
            s_prefix = prefix;
            atomicAdd(&ctl->total_pairs, (unsigned long long)btotal);
            if (tile == n_tiles - 1) {
                const bool over = (unsigned long long)inclusive > capacity;
                ctl->n_sort = over ? (unsigned int)capacity : inclusive;
                ctl->overflow = over ? 1u : 0u;
            }
        }
    }
    __syncthreads();
    const unsigned int prefix = s_prefix;

    // ---- offsets (exclusive scan, global)
    {
        uint4 o;
        o.x = sat_add(prefix, s_scan[tid * 4 + 0]); o.y = sat_add(prefix, s_scan[tid * 4 + 1]);
        o.z = sat_add(prefix, s_scan[tid * 4 + 2]); o.w = sat_add(prefix, s_scan[tid * 4 + 3]);
        *reinterpret_cast<uint4*>(offsets + base + tid * SCAN_ITEMS) = o;
    }
    if (btotal == 0) return;      // nothing visible in this tile of splats (common with tile-row bands): no pairs, no histogram

    // ---- cooperative expansion: output slot j -> (splat, k-th tile of its rectangle)
    int tile_pass[2], n_tile_pass = 0;
    for (int p = 0; p < plan.n_pass; ++p) if (plan.shift[p] >= 32 && n_tile_pass < 2) tile_pass[n_tile_pass++] = p;
    const bool layer = U.mode == SPLAT_MODE_LAYER50;        // second, depth-ordered list
    const bool arrival = U.mode != SPLAT_MODE_TILE;         // LAYER50 / MLAB: the list the pixels walk is in instance order
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
    __syncthreads();
    for (int p = 0; p < FRAME_MAX_PASSES; ++p) {
        const unsigned int v = s_hist[p][tid];
        if (v) atomicAdd(&ctl->hist[p][tid], v);
    }
}

void launch_scan_duplicate(const FrameUniforms& U, const splat_record* rec, const unsigned int* touched,
                           unsigned int* offsets, unsigned long long* keys, unsigned int* vals,
                           unsigned long long* keys_depth, unsigned int* vals_depth,
                           unsigned long long capacity, FrameCtl* ctl, unsigned long long* scan_status,
                           unsigned int epoch, SortPlan plan, cudaStream_t s)
{
    if (U.n == 0) return;
    const unsigned int n_tiles = (U.n + SCAN_TILE_ITEMS - 1) / SCAN_TILE_ITEMS;
    scan_duplicate_kernel<<<n_tiles, SCAN_THREADS, 0, s>>>(U, rec, touched, offsets, keys, vals, keys_depth, vals_depth, capacity, ctl,
                                                           scan_status, epoch, plan, n_tiles);
}
