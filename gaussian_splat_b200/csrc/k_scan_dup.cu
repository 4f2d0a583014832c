// K2: exclusive scan of tiles_touched (tile prefixes come from tile_scan_kernel, so there is no ticket and
// no lookback chain here) fused with (tile | depth) key / value duplication and the radix-sort digit histograms.
// Replaces the hardware tiler's binning of the instanced quads
// (src/instanced_splat_renderer.mm:496-499 draw call).
#include "splat_internal.cuh"

__device__ __forceinline__ unsigned int sat_add(unsigned int a, unsigned int b) {
    const unsigned long long s = (unsigned long long)a + b;
    return s > 0xFFFFFFFFull ? 0xFFFFFFFFu : (unsigned int)s;
}

// What K2 needs of the sort plan: where the two depth digits (bits 0-7, 8-15 of the depth field; absent in MLAB
// frames) and the one or two tile digits (bits 32-39, 40-47 of the key) are counted.
struct DupPlan { int d0, d1, t0, t1; unsigned int tmask0, tmask1; };

// Persistent grid: a CTA walks duplication tiles (1024 splats) with a fixed stride and keeps ONE digit
// histogram for all of them, flushed once at the end (with one CTA per tile, a 6M-splat band frame issued
// 5860 x up to 1024 same-address global atomics). Tiles that produced no pair only write their offsets.
// Per tile: one block scan gives every splat its first pair slot AND the visible splats their rank, so the
// visible ones are compacted into shared memory (first slot, rect, depth field, index) straight from K1's
// 8-byte bin words: no gather from the 48-byte records. The expansion then
// works warp by warp on 32 consecutive pair slots: one binary search for the owner of the first slot, and
// because every visible splat owns at least one slot the other 31 owners lie within the next 32 compact
// entries - each lane finds its own with five shuffles instead of a ten-step search in shared memory.
//
// SPARSE (tile-row bands, where K1 culls up to 97% of the splats): a tile is the BAND_TILE splats of one K1 CTA
// and its entries are the splats K1's visibility bitmap names - 64 words instead of 2048 bin words, and the
// scan / compaction / expansion machinery runs once per ~140 visible splats instead of once per 1024 splats
// (6M-splat band: 31.5M warp instructions at 67% issue before). Entries are taken 1024 at a time, so a tile
// whose splats are all visible still fits the same shared arrays. No per-splat offsets are written (the debug
// fetch derives them from the bin words).
template <bool SPARSE>
__global__ void __launch_bounds__(SCAN_THREADS, 8) scan_duplicate_kernel(
    const __grid_constant__ FrameUniforms U, const unsigned long long* __restrict__ binw,
    unsigned int* __restrict__ offsets,
    unsigned int* __restrict__ keys, unsigned int* __restrict__ vals,
    unsigned int* __restrict__ keys_depth, unsigned int* __restrict__ vals_depth, unsigned long long capacity,
    FrameCtl* ctl, const unsigned int* __restrict__ tile_sums, const unsigned int* __restrict__ scan_prefix,
    const unsigned int* __restrict__ visbits, const DupPlan plan, unsigned int n_tiles, unsigned int n_k1_tiles)
{
    constexpr unsigned int TILE = SPARSE ? BAND_TILE : SCAN_TILE_ITEMS;      // splats per tile
    __shared__ unsigned int s_start[SCAN_TILE_ITEMS + 1];      // first pair slot of the v-th visible splat (+ sentinel)
    __shared__ unsigned int s_rect[SCAN_TILE_ITEMS], s_wdt[SCAN_TILE_ITEMS];   // tx0 | ty0 << 12, rectangle width in tiles
    __shared__ unsigned int s_dfi[SCAN_TILE_ITEMS];            // depth field << 16 | splat index within the tile
    __shared__ unsigned int s_hist[FRAME_MAX_PASSES][256];
    __shared__ unsigned int s_warp[SCAN_THREADS / 32], s_wvis[SCAN_THREADS / 32];
    __shared__ unsigned short s_idx[SPARSE ? BAND_TILE : 1];   // SPARSE: the tile's visible splats, in instance order
    __shared__ unsigned int s_nvis;

    const unsigned int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    pdl_prologue();
    for (int p = 0; p < FRAME_MAX_PASSES; ++p) s_hist[p][tid] = 0;
    const bool layer = U.mode == SPLAT_MODE_LAYER50;        // second, depth-ordered list
    const bool arrival = U.mode != SPLAT_MODE_TILE;         // LAYER50 / MLAB: the list the pixels walk is in instance order
    __syncthreads();

    for (unsigned int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const unsigned int base = tile * TILE;
        constexpr unsigned int CELLS = TILE / (SCAN_TILE_ITEMS / 2);          // K1 pair-count cells per tile
        const unsigned int prefix = __ldg(scan_prefix + tile * (TILE / SCAN_TILE_ITEMS));      // saturated at 2^32 - 1, like the offsets
        unsigned int btotal = 0;
#pragma unroll
        for (unsigned int k = 0; k < CELLS; ++k)
            if (CELLS * tile + k < n_k1_tiles) btotal += __ldg(tile_sums + CELLS * tile + k);
        if (btotal == 0) {      // nothing visible in this tile of splats (common with tile-row bands)
            if (!SPARSE) *reinterpret_cast<uint4*>(offsets + base + tid * SCAN_ITEMS) = make_uint4(prefix, prefix, prefix, prefix);
            continue;
        }
        unsigned int n_ent = SCAN_TILE_ITEMS;       // entries of this tile: every splat, or the visible ones
        if (SPARSE) {
            // visibility bitmap -> list of the visible splats' local indices, in instance order
            if (tid < TILE / 32) {
                unsigned int w = __ldg(visbits + tile * (TILE / 32) + tid);
                const unsigned int c = __popc(w);
                unsigned int incl = c;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) { const unsigned int t = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= d) incl += t; }
                if (lane == 31) s_wvis[warp] = incl;
                __syncwarp();
                // TILE / 32 = 64 words = warps 0 and 1
                asm volatile("bar.sync 1, 64;" ::: "memory");
                unsigned int pos = incl - c + (warp == 1 ? s_wvis[0] : 0u);
                if (tid == TILE / 32 - 1) s_nvis = pos + c;
                while (w) { const unsigned int b = __ffs(w) - 1; s_idx[pos++] = (unsigned short)(tid * 32u + b); w &= w - 1u; }
            }
            __syncthreads();
            n_ent = s_nvis;
        }
        unsigned int carry = 0;         // pairs of the tile's earlier chunks
        for (unsigned int c0 = 0; c0 < n_ent; c0 += SCAN_TILE_ITEMS) {        // one chunk unless SPARSE and > 1024 visible
        // ---- 4 consecutive entries per thread (buffers are padded to the tile size); scan of pairs and of visible splats
        unsigned long long bw[SCAN_ITEMS];
        unsigned int eli[SCAN_ITEMS];                 // the entry's splat index within the tile
        if (SPARSE) {
#pragma unroll
            for (int k = 0; k < SCAN_ITEMS; ++k) {
                const unsigned int e = c0 + tid * SCAN_ITEMS + k;
                eli[k] = e < n_ent ? (unsigned int)s_idx[e] : 0u;
                bw[k] = e < n_ent ? __ldg(binw + base + eli[k]) : 0ull;
            }
        } else {
            const ulonglong2 w01 = *reinterpret_cast<const ulonglong2*>(binw + base + tid * SCAN_ITEMS);
            const ulonglong2 w23 = *reinterpret_cast<const ulonglong2*>(binw + base + tid * SCAN_ITEMS + 2);
            bw[0] = w01.x; bw[1] = w01.y; bw[2] = w23.x; bw[3] = w23.y;
#pragma unroll
            for (int k = 0; k < SCAN_ITEMS; ++k) eli[k] = tid * SCAN_ITEMS + k;
        }
        unsigned int cnt[SCAN_ITEMS];
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; ++k) cnt[k] = binw_count(bw[k]);
        unsigned int tsum = 0, tvis = 0;
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; ++k) { tsum += cnt[k]; tvis += cnt[k] ? 1u : 0u; }
        unsigned int incl = tsum, ivis = tvis;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned int a = __shfl_up_sync(0xFFFFFFFFu, incl, d), b = __shfl_up_sync(0xFFFFFFFFu, ivis, d);
            if (lane >= d) { incl += a; ivis += b; }
        }
        if (lane == 31) { s_warp[warp] = incl; s_wvis[warp] = ivis; }
        __syncthreads();
        unsigned int wbase = 0, vbase = 0, n_vis = 0, ctotal = 0;
#pragma unroll
        for (int k = 0; k < SCAN_THREADS / 32; ++k) {
            const unsigned int a = s_warp[k], b = s_wvis[k];
            if (k < (int)warp) { wbase += a; vbase += b; }
            n_vis += b; ctotal += a;
        }
        unsigned int excl = wbase + incl - tsum;   // first pair slot of this thread's first entry, chunk-local
        unsigned int v = vbase + ivis - tvis;      // rank of this thread's first visible splat

        // ---- offsets; visible splats compacted with their rect / depth field; depth-digit histograms
        uint4 o;
        unsigned int* ov = &o.x;
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; ++k) {
            ov[k] = sat_add(prefix, excl);
            if (cnt[k]) {
                const unsigned int h = (unsigned int)(bw[k] >> 48);
                const unsigned int df = U.order == SPLAT_ORDER_REFERENCE ? (0xFFFFu - h) : h;   // half-depth field (D3)
                s_start[v] = excl;
                s_dfi[v] = (df << 16) | eli[k];              // index < 2048 fits beside the 16-bit depth field
                s_rect[v] = (unsigned int)bw[k] & 0xFFFFFFu;              // tx0 | ty0 << 12
                s_wdt[v] = ((unsigned int)(bw[k] >> 24) & 0xFFFu) - ((unsigned int)bw[k] & 0xFFFu) + 1u;
                if (plan.d0 >= 0) {
                    atomicAdd(&s_hist[plan.d0][df & 0xFFu], cnt[k]);
                    atomicAdd(&s_hist[plan.d1][df >> 8], cnt[k]);
                }
                ++v;
            }
            excl += cnt[k];
        }
        if (!SPARSE) *reinterpret_cast<uint4*>(offsets + base + tid * SCAN_ITEMS) = o;
        if (tid == 0) s_start[n_vis] = ctotal;              // sentinel: end of the last owner
        __syncthreads();

        // ---- expansion: a warp takes 32 consecutive pair slots at a time
        for (unsigned int j0 = warp * 32u; j0 < ctotal; j0 += SCAN_THREADS) {
            const unsigned int j = j0 + lane;
            // owner of slot j0: last compact entry whose first slot is <= j0 (uniform across the warp)
            unsigned int lo = 0, hi = n_vis;
            while (hi - lo > 1) { const unsigned int mid = (lo + hi) >> 1; if (s_start[mid] <= j0) lo = mid; else hi = mid; }
            // ends of the 32 owners from there on; lane's owner = lo + #{ends <= j}
            const unsigned int end_l = s_start[min(lo + lane + 1u, n_vis)];
            unsigned int c = 0;
#pragma unroll
            for (int st = 16; st > 0; st >>= 1) {
                const unsigned int e = __shfl_sync(0xFFFFFFFFu, end_l, c + st - 1);
                if (e <= j) c += st;
            }
            if (j < ctotal) {
                const unsigned int vi = lo + c;
                const unsigned int k = j - s_start[vi];
                const unsigned int rect = s_rect[vi], wdt = s_wdt[vi], dfi = s_dfi[vi];
                const unsigned int tx0 = rect & 0xFFFu, ty0 = rect >> 12;
                const unsigned int ry = k / wdt, rx = k - ry * wdt;
                const unsigned int t = (ty0 + ry) * (unsigned int)U.tiles_x + (tx0 + rx);
                const unsigned int splat = base + (dfi & 0xFFFFu), df = dfi >> 16;
{   // tile-digit histograms: ONE match on the whole tile id serves both digits, the lowest peer adds the
                    // group's size to each (a match per digit cost twice the ADU time: K2 187 -> 152 us at 6M / 4K).
                    // A warp's 32 consecutive slots belong to about eight splats at unrelated screen positions;
                    // tried instead of the match: a plain shared atomic per lane (40% slower: two LSU cycles per
                    // lane, serialised on equal addresses) and a shuffle + ballot loop over the distinct values
                    // of the high digit (60% slower)
                    const unsigned int peers = __match_any_sync(__activemask(), t);
                    if ((peers & ((1u << lane) - 1u)) == 0) {
                        const unsigned int c = __popc(peers);
                        atomicAdd(&s_hist[plan.t0][t & plan.tmask0], c);
                        if (plan.t1 >= 0) atomicAdd(&s_hist[plan.t1][(t >> 8) & plan.tmask1], c);
                    }
                }
                const unsigned long long g = (unsigned long long)prefix + carry + j;
                if (g < capacity) {
                    // TILE: tile | depth. LAYER50 / MLAB: tile alone (slots are in instance order and the sort is
                    // stable); LAYER50 adds a second copy in tile | depth order for the composite.
                    keys[g] = (t << 16) | (arrival ? 0u : df);
                    vals[g] = splat;
                    if (layer) {
                        keys_depth[g] = (t << 16) | df;
                        vals_depth[g] = splat;
                    }
                }
            }
        }
        __syncthreads();            // the compact arrays are reused by the next chunk / tile
        carry += ctotal;
        }
    }
    __syncthreads();
    for (int p = 0; p < FRAME_MAX_PASSES; ++p) {
        const unsigned int v = s_hist[p][tid];
        if (v) atomicAdd(&ctl->hist[p][tid], v);
    }
}

void launch_scan_duplicate(const FrameUniforms& U, const unsigned long long* binw,
                           unsigned int* offsets, unsigned int* keys, unsigned int* vals,
                           unsigned int* keys_depth, unsigned int* vals_depth,
                           unsigned long long capacity, FrameCtl* ctl, const unsigned int* tile_sums,
                           const unsigned int* scan_prefix, const unsigned int* visbits, SortPlan plan, int sm_count, cudaStream_t s)
{
    if (U.n == 0) return;
    DupPlan dp{-1, -1, -1, -1, 0u, 0u};
    for (int p = 0; p < plan.n_pass; ++p) {
        if (plan.shift[p] == 0) dp.d0 = plan.slot[p];
        else if (plan.shift[p] == 8) dp.d1 = plan.slot[p];
        else if (plan.shift[p] == 16) { dp.t0 = plan.slot[p]; dp.tmask0 = plan.mask[p]; }
        else if (plan.shift[p] == 24) { dp.t1 = plan.slot[p]; dp.tmask1 = plan.mask[p]; }
    }
    if (dp.t0 < 0) { dp.t0 = FRAME_MAX_PASSES - 1; dp.tmask0 = 0; }      // single-tile frame: no tile digit is sorted; count into an unused row
    const unsigned int n_k1 = (U.n + SCAN_TILE_ITEMS / 2 - 1) / (SCAN_TILE_ITEMS / 2);
    const bool sparse = frame_is_band(U);
    const unsigned int per = sparse ? BAND_TILE : SCAN_TILE_ITEMS;
    const unsigned int n_tiles = (U.n + per - 1) / per;
    const unsigned int grid = n_tiles < (unsigned int)sm_count * 8u ? n_tiles : (unsigned int)sm_count * 8u;
    if (sparse)
        launch_pdl(scan_duplicate_kernel<true>, dim3(grid), dim3(SCAN_THREADS), 0, s, U, binw, offsets, keys, vals, keys_depth, vals_depth, capacity, ctl,
                   tile_sums, scan_prefix, visbits, dp, n_tiles, n_k1);
    else
        launch_pdl(scan_duplicate_kernel<false>, dim3(grid), dim3(SCAN_THREADS), 0, s, U, binw, offsets, keys, vals, keys_depth, vals_depth, capacity, ctl,
                   tile_sums, scan_prefix, visbits, dp, n_tiles, n_k1);
}
