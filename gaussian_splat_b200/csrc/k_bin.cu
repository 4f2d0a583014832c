// K2: binning of the visible splats into per-tile lists. Replaces the hardware tiler's binning of the instanced
// quads (src/instanced_splat_renderer.mm:496-499 draw call) and the per-pixel insertion sorts of
// shaders/gaussian_splat_tile.metal:240-249 / gaussian_splat_50layer.metal:198-206.
//
//   bin_pairs_kernel     every (splat, tile of its rectangle) candidate is tested (D9' tile_touch); a passing pair draws
//                        its slot in the tile's list from the tile's counter (one L2 atomic) and is parked at its
//                        candidate position in (tmp_key, tmp_ts)
//   tile_offsets_kernel  exclusive scan of the T tile counters -> bounds[T + 1], D, the longest list
//   scatter_pairs_kernel pairs[bounds[tile] + slot] = depth field << 32 | instance index
//   tile_sort_kernel     one CTA per tile sorts its list on the whole 64-bit pair. Instance indices are unique within a
//                        list, so the key is total and the result is what a stable sort of instance-ordered pairs on
//                        tile | depth gives, whatever order the atomics handed the slots out in.
// Round 1 sorted the frame's (tile << 16 | depth) keys with four onesweep passes over all D pairs and searched the tile
// boundaries afterwards: at frame sizes (1.6M pairs = less than one wave of sort tiles) that was four latency-bound
// chains of load / rank / lookback / scatter, 66 + 10 us of a 346 us frame at 23% of HBM. Here every pair is written
// twice and read twice, and the sort runs on T independent lists of a few hundred pairs.
#include "splat_math.cuh"

#define HOLE 0xFFFFFFFFFFFFFFFFull

// Persistent grid: a CTA walks tiles of 1024 splats with a fixed stride. Per tile: one block scan gives every splat its
// first candidate slot AND the visible splats their rank, so the visible ones are compacted into shared memory (first
// slot, rectangle, depth field, index) straight from K1's 8-byte bin words. The expansion then works warp by warp on 32
// consecutive candidate slots: one binary search for the owner of the first slot, and because every visible splat owns
// at least one slot the other 31 owners lie within the next 32 compact entries - each lane finds its own with five
// shuffles. The lanes of one splat read the same record row (one L2 request), run the tile test and draw a slot.
__global__ void __launch_bounds__(SCAN_THREADS, 8) bin_pairs_kernel(
    const __grid_constant__ FrameUniforms U, const unsigned long long* binw, const splat_record_row* rec,
    unsigned long long* __restrict__ tmp_key, unsigned long long* __restrict__ tmp_ts,
    unsigned long long capacity, FrameCtl* ctl, const unsigned int* tile_sums, const unsigned int* scan_prefix,
    unsigned int n_tiles, unsigned int n_cells)
{
    __shared__ unsigned int s_start[SCAN_TILE_ITEMS + 1];      // first candidate slot of the v-th visible splat (+ sentinel)
    __shared__ unsigned int s_rect[SCAN_TILE_ITEMS], s_wdt[SCAN_TILE_ITEMS];   // tx0 | ty0 << 12, rectangle width in tiles
    __shared__ unsigned int s_dfi[SCAN_TILE_ITEMS];            // depth field << 16 | splat index within the tile
    __shared__ unsigned int s_warp[SCAN_THREADS / 32], s_wvis[SCAN_THREADS / 32];

    const unsigned int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    pdl_prologue();
    if (__ldcg(&ctl->overflow)) return;
    unsigned int* const tile_count = ctl_tile_count(ctl);
    const bool by_instance = U.mode == SPLAT_MODE_MLAB;       // the only list MLAB walks is in instance order: no depth in the key

    for (unsigned int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const unsigned int base = tile * SCAN_TILE_ITEMS;
        unsigned int btotal = __ldcg(tile_sums + 2 * tile);
        if (2 * tile + 1 < n_cells) btotal += __ldcg(tile_sums + 2 * tile + 1);
        if (btotal == 0) continue;      // nothing visible in this tile of splats
        const unsigned int prefix = __ldcg(scan_prefix + tile);
        // ---- 4 consecutive splats per thread (buffers are padded to the tile size); scan of candidates and of visible splats
        unsigned long long bw[SCAN_ITEMS];
        {
            const ulonglong2 w01 = __ldcg(reinterpret_cast<const ulonglong2*>(binw + base + tid * SCAN_ITEMS));
            const ulonglong2 w23 = __ldcg(reinterpret_cast<const ulonglong2*>(binw + base + tid * SCAN_ITEMS + 2));
            bw[0] = w01.x; bw[1] = w01.y; bw[2] = w23.x; bw[3] = w23.y;
        }
        unsigned int cnt[SCAN_ITEMS];
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; ++k) cnt[k] = base + tid * SCAN_ITEMS + k < U.n ? binw_count(bw[k]) : 0u;
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
        unsigned int excl = wbase + incl - tsum;   // first candidate slot of this thread's first splat, tile-local
        unsigned int v = vbase + ivis - tvis;      // rank of this thread's first visible splat
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; ++k) {
            if (cnt[k]) {
                const unsigned int h = (unsigned int)(bw[k] >> 48);
                const unsigned int df = by_instance ? 0u : (U.order == SPLAT_ORDER_REFERENCE ? (0xFFFFu - h) : h);   // half-depth field (D3)
                s_start[v] = excl;
                s_dfi[v] = (df << 16) | (tid * SCAN_ITEMS + k);
                s_rect[v] = (unsigned int)bw[k] & 0xFFFFFFu;              // tx0 | ty0 << 12
                s_wdt[v] = ((unsigned int)(bw[k] >> 24) & 0xFFFu) - ((unsigned int)bw[k] & 0xFFFu) + 1u;
                ++v;
            }
            excl += cnt[k];
        }
        if (tid == 0) s_start[n_vis] = ctotal;              // sentinel: end of the last owner
        __syncthreads();

        // ---- expansion: a warp takes 32 consecutive candidate slots at a time
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
            const unsigned long long g = (unsigned long long)prefix + j;
            if (j < ctotal && g < capacity) {
                const unsigned int vi = lo + c;
                const unsigned int k = j - s_start[vi];
                const unsigned int rect = s_rect[vi], wdt = s_wdt[vi], dfi = s_dfi[vi];
                const unsigned int tx0 = rect & 0xFFFu, ty0 = rect >> 12;
                const unsigned int ry = k / wdt, rx = k - ry * wdt;
                const unsigned int tx = tx0 + rx, ty = ty0 + ry;
                const unsigned int splat = base + (dfi & 0xFFFFu), df = dfi >> 16;
                const float4* r = reinterpret_cast<const float4*>(rec + splat);
                const float4 A = __ldcg(r), B = __ldcg(r + 1), R = __ldcg(r + 3);
                if (tile_touch(A.x, A.y, A.z, A.w, B.x, B.y, R.x, R.y, (int)tx, (int)ty, U.w, U.h)) {
                    const unsigned int t = ty * (unsigned int)U.tiles_x + tx;
                    const unsigned int slot = atomicAdd(tile_count + t, 1u);
                    tmp_key[g] = ((unsigned long long)df << 32) | splat;
                    tmp_ts[g] = ((unsigned long long)t << 32) | slot;
                } else tmp_ts[g] = HOLE;
            }
        }
        __syncthreads();            // the compact arrays are reused by the next tile
    }
}

void launch_bin_pairs(const FrameUniforms& U, const unsigned long long* binw, const splat_record_row* rec,
                      unsigned long long* tmp_key, unsigned long long* tmp_ts, unsigned long long capacity, FrameCtl* ctl,
                      const unsigned int* tile_sums, const unsigned int* scan_prefix, int sm_count, cudaStream_t s)
{
    if (U.n == 0) return;
    const unsigned int n_cells = (U.n + SCAN_TILE_ITEMS / 2 - 1) / (SCAN_TILE_ITEMS / 2);
    const unsigned int n_tiles = (U.n + SCAN_TILE_ITEMS - 1) / SCAN_TILE_ITEMS;
    const unsigned int grid = n_tiles < (unsigned int)sm_count * 8u ? n_tiles : (unsigned int)sm_count * 8u;
    launch_pdl(bin_pairs_kernel, dim3(grid), dim3(SCAN_THREADS), 0, s, U, binw, rec, tmp_key, tmp_ts, capacity, ctl,
               tile_sums, scan_prefix, n_tiles, n_cells);
}

// bounds[t] = pairs of tiles 0 .. t-1 (t = 0 .. T); one CTA, a contiguous chunk of counters per thread.
struct FrameStatusSlotDev { unsigned long long total_pairs, total_cand; unsigned int overflow, max_tile_pairs, frame_no, pad; };
__global__ void __launch_bounds__(1024) tile_offsets_kernel(FrameCtl* ctl, unsigned int n_tiles, unsigned int* __restrict__ bounds,
                                                            volatile FrameStatusSlotDev* slot, unsigned int frame_no)
{
    __shared__ unsigned int s_wsum[32], s_wmax[32];
    pdl_prologue();
    const unsigned int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned int* tile_count = ctl_tile_count(ctl);
    const bool over = __ldcg(&ctl->overflow) != 0;
    const unsigned int per = (n_tiles + 1023u) / 1024u;
    const unsigned int j0 = min(tid * per, n_tiles), j1 = min(j0 + per, n_tiles);
    unsigned int part = 0, big = 0;
    for (unsigned int j = j0; j < j1; ++j) { const unsigned int c = over ? 0u : __ldcg(tile_count + j); part += c; big = max(big, c); }
    unsigned int incl = part;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const unsigned int x = __shfl_up_sync(0xFFFFFFFFu, incl, d);
        if (lane >= (unsigned)d) incl += x;
        big = max(big, __shfl_xor_sync(0xFFFFFFFFu, big, d));
    }
    if (lane == 31) s_wsum[warp] = incl;
    if (lane == 0) s_wmax[warp] = big;
    __syncthreads();
    unsigned int run = incl - part, total = 0, mx = 0;
    for (int k = 0; k < 32; ++k) { const unsigned int x = s_wsum[k]; if (k < (int)warp) run += x; total += x; mx = max(mx, s_wmax[k]); }
    for (unsigned int j = j0; j < j1; ++j) { bounds[j] = run; run += over ? 0u : __ldcg(tile_count + j); }
    if (tid == 0) {
        bounds[n_tiles] = total;
        const unsigned long long cand = __ldcg(&ctl->total_cand);
        ctl->total_pairs = over ? cand : (unsigned long long)total;
        ctl->max_tile_pairs = mx;
        // the frame's status, straight into page-locked host memory; the frame number is written last
        slot->total_pairs = over ? cand : (unsigned long long)total; slot->total_cand = cand;
        slot->overflow = over ? 1u : 0u; slot->max_tile_pairs = mx;
        __threadfence_system();
        slot->frame_no = frame_no;
    }
}

void launch_tile_offsets(const FrameUniforms& U, FrameCtl* ctl, unsigned int* bounds, void* status_slot, unsigned int frame_no, cudaStream_t s)
{
    launch_pdl(tile_offsets_kernel, dim3(1), dim3(1024), 0, s, ctl, (unsigned int)(U.tiles_x * U.tiles_y), bounds,
               (volatile FrameStatusSlotDev*)status_slot, frame_no);
}

__global__ void __launch_bounds__(256) scatter_pairs_kernel(const FrameCtl* ctl, const unsigned long long* tmp_key,
                                                            const unsigned long long* tmp_ts, const unsigned int* bounds,
                                                            unsigned long long* __restrict__ pairs)
{
    pdl_prologue();
    if (__ldcg(&ctl->overflow)) return;
    const unsigned int n = __ldcg(&ctl->n_cand);
    for (unsigned int g = blockIdx.x * 256u + threadIdx.x; g < n; g += gridDim.x * 256u) {
        const unsigned long long ts = __ldcg(tmp_ts + g);
        if (ts == HOLE) continue;
        pairs[__ldcg(bounds + (unsigned int)(ts >> 32)) + (unsigned int)ts] = __ldcg(tmp_key + g);
    }
}

void launch_scatter_pairs(const FrameCtl* ctl, const unsigned long long* tmp_key, const unsigned long long* tmp_ts,
                          const unsigned int* bounds, unsigned long long* pairs, int sm_count, cudaStream_t s)
{
    launch_pdl(scatter_pairs_kernel, dim3(sm_count * 8), dim3(256), 0, s, ctl, tmp_key, tmp_ts, bounds, pairs);
}

// ------------------------------------------------------------------ per-tile sort
#define TS_THREADS 128
#define TS_MAX_E 16                                   // pairs per thread: lists of up to 2048 pairs are sorted in registers
#define TS_CHUNK (TS_THREADS * TS_MAX_E)              // pairs the shared buffer holds
#define SENTINEL 0xFFFFFFFFFFFFFFFFull

// Bitonic network over P = 128 E keys, key i = e * 128 + tid held by thread tid in x[e]: partners at distance >= 128 sit
// in the same thread (registers), at distance 32 and 64 in another warp (exchanged through shared memory), below that in
// the same warp (shuffles). Position i keeps the smaller key when bit j of i is clear in an ascending block.
__device__ __forceinline__ unsigned long long shfl_xor64(unsigned long long v, int m) {
    const unsigned int lo = __shfl_xor_sync(0xFFFFFFFFu, (unsigned int)v, m), hi = __shfl_xor_sync(0xFFFFFFFFu, (unsigned int)(v >> 32), m);
    return ((unsigned long long)hi << 32) | lo;
}
template <int E>
__device__ __forceinline__ void exchange_stage(unsigned long long (&x)[E], unsigned long long* s_x, unsigned int tid, int j, int k) {
    // partner in another warp: through shared memory
#pragma unroll
    for (int e = 0; e < E; ++e) s_x[e * TS_THREADS + tid] = x[e];
    __syncthreads();
#pragma unroll
    for (int e = 0; e < E; ++e) {
        const unsigned int i = e * TS_THREADS + tid;
        const unsigned long long p = s_x[i ^ j];
        const bool keep_min = ((i & j) == 0) == ((i & k) == 0);
        x[e] = (keep_min == (x[e] < p)) ? x[e] : p;
    }
    __syncthreads();
}
template <int E>
__device__ __forceinline__ void shuffle_stage(unsigned long long (&x)[E], unsigned int tid, int j, int k) {
#pragma unroll
    for (int e = 0; e < E; ++e) {
        const unsigned int i = e * TS_THREADS + tid;
        const unsigned long long p = shfl_xor64(x[e], j);
        const bool keep_min = ((i & j) == 0) == ((i & k) == 0);
        x[e] = (keep_min == (x[e] < p)) ? x[e] : p;
    }
}
template <int E, int JE>
__device__ __forceinline__ void register_stage(unsigned long long (&x)[E], int k) {
    // partner e ^ JE in the same thread; bits of i at and above 128 come from e alone
#pragma unroll
    for (int e = 0; e < E; ++e) {
        if ((e & JE) == 0) {
            const bool asc = ((e * TS_THREADS) & k) == 0;
            const unsigned long long a = x[e], b = x[e | JE];
            const bool sw = (a > b) == asc;
            x[e] = sw ? b : a; x[e | JE] = sw ? a : b;
        }
    }
}
template <int E, int JE>
struct RegisterStages {
    static __device__ __forceinline__ void run(unsigned long long (&x)[E], int k) {
        register_stage<E, JE>(x, k);
        RegisterStages<E, JE / 2>::run(x, k);
    }
};
template <int E>
struct RegisterStages<E, 0> { static __device__ __forceinline__ void run(unsigned long long (&)[E], int) {} };

template <int E, int M>     // level k = 128 << M (M = 1 .. log2 E): register stages 64 << M .. 128, then the tail
struct UpperLevels {
    static __device__ __forceinline__ void run(unsigned long long (&x)[E], unsigned long long* s_x, unsigned int tid) {
        UpperLevels<E, M - 1>::run(x, s_x, tid);
        constexpr int k = TS_THREADS << M;
        if constexpr (k <= E * TS_THREADS) {
            RegisterStages<E, (k / 2) / TS_THREADS>::run(x, k);
            exchange_stage<E>(x, s_x, tid, 64, k);
            exchange_stage<E>(x, s_x, tid, 32, k);
            for (int j = 16; j >= 1; j >>= 1) shuffle_stage<E>(x, tid, j, k);
        }
    }
};
template <int E>
struct UpperLevels<E, 0> { static __device__ __forceinline__ void run(unsigned long long (&)[E], unsigned long long*, unsigned int) {} };

template <int E>
__device__ __forceinline__ void bitonic_sort_regs(unsigned long long (&x)[E], unsigned long long* s_x, unsigned int tid) {
    for (int k = 2; k <= 32; k <<= 1)
        for (int j = k >> 1; j >= 1; j >>= 1) shuffle_stage<E>(x, tid, j, k);
    exchange_stage<E>(x, s_x, tid, 32, 64);
    for (int j = 16; j >= 1; j >>= 1) shuffle_stage<E>(x, tid, j, 64);
    exchange_stage<E>(x, s_x, tid, 64, 128);
    exchange_stage<E>(x, s_x, tid, 32, 128);
    for (int j = 16; j >= 1; j >>= 1) shuffle_stage<E>(x, tid, j, 128);
    UpperLevels<E, 4>::run(x, s_x, tid);        // levels 256 .. 2048, those beyond 128 E compiled out
}

template <int E>
__device__ __forceinline__ void sort_list_regs(unsigned long long* seg, unsigned int* arrival, unsigned int len, bool layer,
                                               unsigned long long* s_x, unsigned int tid) {
    unsigned long long x[E];
#pragma unroll
    for (int e = 0; e < E; ++e) { const unsigned int i = e * TS_THREADS + tid; x[e] = i < len ? __ldcg(seg + i) : SENTINEL; }
    bitonic_sort_regs<E>(x, s_x, tid);
#pragma unroll
    for (int e = 0; e < E; ++e) { const unsigned int i = e * TS_THREADS + tid; if (i < len) seg[i] = x[e]; }
    if (!layer) return;
    // LAYER50: the same pairs once more in instance (= arrival) order
#pragma unroll
    for (int e = 0; e < E; ++e) if (x[e] != SENTINEL) x[e] &= 0xFFFFFFFFull;
    bitonic_sort_regs<E>(x, s_x, tid);
#pragma unroll
    for (int e = 0; e < E; ++e) { const unsigned int i = e * TS_THREADS + tid; if (i < len) arrival[i] = (unsigned int)x[e]; }
}

// Lists longer than TS_CHUNK: the same network run by the one CTA on the list in global memory, every comparator
// ascending (merging a block starts by comparing i with its mirror image in the block), so positions beyond the end of
// the list act as +infinity without being touched. Strides of TS_CHUNK and more go through global memory, the rest of a
// level runs on TS_CHUNK pairs at a time in shared memory. Slow (one CTA, O(n log^2 n)) but exact, and rare: the
// benchmark scenes' longest lists are 800 .. 1600 pairs.
__device__ __forceinline__ void ce_mem(unsigned long long* a, unsigned long long* b, unsigned long long mask) {
    const unsigned long long x = *a, y = *b;
    if ((x & mask) > (y & mask)) { *a = y; *b = x; }
}
__device__ __forceinline__ void ce_mem_l2(unsigned long long* a, unsigned long long* b, unsigned long long mask) {      // global memory, L2 only
    const unsigned long long x = __ldcg(a), y = __ldcg(b);
    if ((x & mask) > (y & mask)) { __stcg(a, y); __stcg(b, x); }
}
__device__ void sort_list_global(unsigned long long* seg, unsigned int len, unsigned long long mask, unsigned long long* s_x, unsigned int tid) {
    unsigned int P = TS_CHUNK;
    while (P < len) P <<= 1;
    for (unsigned int k = 2; k <= P; k <<= 1) {
        bool mirror = true;
        for (unsigned int j = k >> 1; j >= 1; j >>= 1, mirror = false) {
            if (k > TS_CHUNK && (mirror || j >= TS_CHUNK)) {
                for (unsigned int p = tid; p < P / 2; p += TS_THREADS) {
                    unsigned int lo, hi;
                    if (mirror) { const unsigned int blk = p / (k >> 1), off = p % (k >> 1); lo = blk * k + off; hi = blk * k + (k - 1u - off); }
                    else { lo = (p / j) * (2u * j) + (p % j); hi = lo + j; }
                    if (hi < len) ce_mem_l2(seg + lo, seg + hi, mask);
                }
                __syncthreads();
                continue;
            }
            // the remaining stages of this level stay inside aligned TS_CHUNK blocks: run them in shared memory
            for (unsigned int c0 = 0; c0 < len; c0 += TS_CHUNK) {
                for (unsigned int i = tid; i < TS_CHUNK; i += TS_THREADS) s_x[i] = c0 + i < len ? __ldcg(seg + c0 + i) : SENTINEL;
                __syncthreads();
                bool mir = mirror;
                for (unsigned int jj = j; jj >= 1; jj >>= 1, mir = false) {
                    for (unsigned int p = tid; p < TS_CHUNK / 2; p += TS_THREADS) {
                        unsigned int lo, hi;
                        if (mir) { const unsigned int blk = p / (k >> 1), off = p % (k >> 1); lo = blk * k + off; hi = blk * k + (k - 1u - off); }
                        else { lo = (p / jj) * (2u * jj) + (p % jj); hi = lo + jj; }
                        ce_mem(s_x + lo, s_x + hi, mask);
                    }
                    __syncthreads();
                }
                for (unsigned int i = tid; i < TS_CHUNK; i += TS_THREADS) if (c0 + i < len) __stcg(seg + c0 + i, s_x[i]);
                __syncthreads();
            }
            break;
        }
    }
}

__global__ void __launch_bounds__(TS_THREADS) tile_sort_kernel(const __grid_constant__ FrameUniforms U, const FrameCtl* ctl,
                                                               const unsigned int* bounds, unsigned long long* pairs,
                                                               unsigned int* arrival)
{
    __shared__ unsigned long long s_x[TS_CHUNK];
    pdl_prologue();
    const unsigned int tile = (unsigned int)((U.row_begin + blockIdx.y) * U.tiles_x) + blockIdx.x;
    const unsigned int b0 = __ldcg(bounds + tile), len = __ldcg(bounds + tile + 1) - b0;
    const unsigned int tid = threadIdx.x;
    if (len <= 1 && U.mode != SPLAT_MODE_LAYER50) return;
    if (len == 0) return;
    unsigned long long* seg = pairs + b0;
    const bool layer = U.mode == SPLAT_MODE_LAYER50;
    if (len <= 32) {
        // one warp, shuffles only
        if (tid >= 32) return;
        unsigned long long x = tid < len ? __ldcg(seg + tid) : SENTINEL;
        for (int pass = 0; pass < (layer ? 2 : 1); ++pass) {
            for (int k = 2; k <= 32; k <<= 1)
                for (int j = k >> 1; j >= 1; j >>= 1) {
                    const unsigned long long p = shfl_xor64(x, j);
                    const bool keep_min = ((tid & j) == 0) == ((tid & k) == 0);
                    x = (keep_min == (x < p)) ? x : p;
                }
            if (pass == 0) { if (tid < len) seg[tid] = x; if (x != SENTINEL) x &= 0xFFFFFFFFull; }
            else if (tid < len) arrival[b0 + tid] = (unsigned int)x;
        }
        return;
    }
    if (len <= 1 * TS_THREADS) sort_list_regs<1>(seg, arrival + b0, len, layer, s_x, tid);
    else if (len <= 2 * TS_THREADS) sort_list_regs<2>(seg, arrival + b0, len, layer, s_x, tid);
    else if (len <= 4 * TS_THREADS) sort_list_regs<4>(seg, arrival + b0, len, layer, s_x, tid);
    else if (len <= 8 * TS_THREADS) sort_list_regs<8>(seg, arrival + b0, len, layer, s_x, tid);
    else if (len <= 16 * TS_THREADS) sort_list_regs<16>(seg, arrival + b0, len, layer, s_x, tid);
    else {
        if (layer) {
            // instance order first (on the low words), copied out, then the depth order in place
            sort_list_global(seg, len, 0xFFFFFFFFull, s_x, tid);
            for (unsigned int i = tid; i < len; i += TS_THREADS) arrival[b0 + i] = (unsigned int)__ldcg(seg + i);
            __syncthreads();
        }
        sort_list_global(seg, len, ~0ull, s_x, tid);
    }
}

void launch_tile_sort(const FrameUniforms& U, const FrameCtl* ctl, const unsigned int* bounds, unsigned long long* pairs,
                      unsigned int* arrival, cudaStream_t s)
{
    const int rows = U.row_end - U.row_begin;
    if (rows <= 0 || U.tiles_x <= 0) return;
    launch_pdl(tile_sort_kernel, dim3(U.tiles_x, rows), dim3(TS_THREADS), 0, s, U, ctl, bounds, pairs, arrival);
}
