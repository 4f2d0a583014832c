// K2: from K1's parked (splat, tile) pairs to sorted per-tile lists. Replaces the hardware tiler's binning of the
// instanced quads (src/instanced_splat_renderer.mm:496-499 draw call) and the per-pixel insertion sorts of
// shaders/gaussian_splat_tile.metal:240-249 / gaussian_splat_50layer.metal:198-206.
//
//   (K1)                 every (splat, tile of its rectangle) candidate that passes the D9' tile test draws its slot from its
//                        tile's counter and is parked (key = depth field << 32 | instance index; tile, slot) - k_preprocess.cu
//   tile_offsets_kernel  exclusive scan of the T tile counters -> bounds[T + 1], D,
//                        the longest list
//   scatter_pairs_kernel pairs[bounds[tile] + slot] = key
//   tile_sort_kernel     one CTA per tile sorts its list on the whole 64-bit pair. Instance indices are unique within a
//                        list, so the key is total and the result is what a stable sort of instance-ordered pairs on
//                        tile | depth gives, whatever order the atomics handed the slots out in.
// Round 1 sorted the frame's (tile << 16 | depth) keys with four onesweep passes over all D pairs and searched the tile
// boundaries afterwards: at frame sizes (1.6M pairs = less than one wave of sort tiles) that was four latency-bound
// chains of load / rank / lookback / scatter, 66 + 10 us of a 346 us frame at 23% of HBM. Here every pair is written
// twice and read twice, and the sort runs on T independent lists of a few hundred pairs.
#include "splat_math.cuh"

// bounds[t] = first pair of tile t's list, for the tiles [first, first + count] of the frame's band (the blend and the sort
// read nothing else). One CTA: a thread owns a contiguous run of counters, read with up to V independent 16-byte loads (one
// L2 round trip), scanned in registers, written back as vectors. The counter and bounds arrays are padded to
// 65536 + 8 words; the run starts at `first` rounded down to a vector boundary, counters outside the band count as 0.
template <int V>      // 16-byte vectors per thread: bands of up to 4096 V - 4 tiles
__global__ void __launch_bounds__(1024) tile_offsets_kernel(FrameCtl* ctl, unsigned int first, unsigned int count,
                                                            unsigned int* __restrict__ bounds)
{
    __shared__ unsigned int s_wsum[32], s_wmax[32];
    pdl_prologue();
    const unsigned int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned int* tile_count = ctl_tile_count(ctl);
    const bool over = __ldcg(&ctl->overflow) != 0;
    const unsigned int lo = first & ~3u, end = first + count;                // counters [lo, lo + 4096 V) are looked at
    const unsigned int per4 = (end - lo + 4095u) / 4096u;                    // 16-byte vectors per thread (<= V)
    const unsigned int mine = lo + tid * per4 * 4u;
    uint4 v[V];
#pragma unroll
    for (int q = 0; q < V; ++q)
        if ((unsigned)q < per4 && mine + 4u * q < end) v[q] = __ldcg(reinterpret_cast<const uint4*>(tile_count + mine) + q);
    unsigned int part = 0, big = 0;
#pragma unroll
    for (int q = 0; q < V; ++q)
        if ((unsigned)q < per4) {
            unsigned int* e = &v[q].x;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const unsigned int t = mine + 4u * q + k;
                if (over || t < first || t >= end) e[k] = 0u;
                part += e[k]; big = max(big, e[k]);
            }
        }
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
#pragma unroll
    for (int k = 0; k < 32; ++k) { const unsigned int x = s_wsum[k]; if (k < (int)warp) run += x; total += x; mx = max(mx, s_wmax[k]); }
#pragma unroll
    for (int q = 0; q < V; ++q)
        if ((unsigned)q < per4 && mine + 4u * q <= end) {        // <= : the vector holding bounds[end] = total is written too
            uint4 o;
            o.x = run; run += v[q].x; o.y = run; run += v[q].y; o.z = run; run += v[q].z; o.w = run; run += v[q].w;
            reinterpret_cast<uint4*>(bounds + mine)[q] = o;
        }
    if (tid == 0) {
        bounds[end] = total;
        ctl->total_pairs = total;
        ctl->max_tile_pairs = mx;
    }
}

void launch_tile_offsets(const FrameUniforms& U, FrameCtl* ctl, unsigned int* bounds, cudaStream_t s)
{
    const unsigned int first = (unsigned int)(U.row_begin * U.tiles_x), count = (unsigned int)(max(U.row_end - U.row_begin, 0) * U.tiles_x);
    const unsigned int span = first + count - (first & ~3u) + 4u;
    if (span <= 16384u) launch_pdl(tile_offsets_kernel<4>, dim3(1), dim3(1024), 0, s, ctl, first, count, bounds);
    else if (span <= 32768u) launch_pdl(tile_offsets_kernel<8>, dim3(1), dim3(1024), 0, s, ctl, first, count, bounds);
    else launch_pdl(tile_offsets_kernel<16>, dim3(1), dim3(1024), 0, s, ctl, first, count, bounds);
}

#define SCATTER_WAYS 8
// A CTA takes eight parking chunks at a time, one slot per thread and chunk: the (key; tile, slot) loads, the bounds lookups
// and the stores are eight independent chains, no atomics (K1 drew the slots). Thread 0 of CTA 0 also leaves the frame's
// status in page-locked host memory (the frame number last), where the next entry-point call finds it without a copy.
struct FrameStatusSlotDev { unsigned long long total_pairs, total_slots; unsigned int overflow, max_tile_pairs, frame_no, pad; };
__global__ void __launch_bounds__(PARK_CHUNK) scatter_pairs_kernel(const FrameCtl* ctl, const unsigned long long* park_key,
                                                                   const unsigned long long* park_ts, const unsigned int* chunk_fill,
                                                                   const unsigned int* bounds, unsigned long long* __restrict__ pairs,
                                                                   volatile FrameStatusSlotDev* slot, unsigned int frame_no)
{
    pdl_prologue();
    const unsigned int over = __ldcg(&ctl->overflow), n_chunks = __ldcg(&ctl->n_chunks);
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        slot->total_pairs = __ldcg(&ctl->total_pairs); slot->total_slots = (unsigned long long)n_chunks * PARK_CHUNK;
        slot->overflow = over; slot->max_tile_pairs = __ldcg(&ctl->max_tile_pairs);
        __threadfence_system();
        slot->frame_no = frame_no;
    }
    if (over) return;
    const unsigned int tid = threadIdx.x;
    for (unsigned int c0 = blockIdx.x * SCATTER_WAYS; c0 < n_chunks; c0 += gridDim.x * SCATTER_WAYS) {
        unsigned long long ts[SCATTER_WAYS], k[SCATTER_WAYS]; bool on[SCATTER_WAYS];
#pragma unroll
        for (int q = 0; q < SCATTER_WAYS; ++q) on[q] = c0 + q < n_chunks && tid < __ldcg(chunk_fill + c0 + q);
#pragma unroll
        for (int q = 0; q < SCATTER_WAYS; ++q)
            if (on[q]) { const size_t g = (size_t)(c0 + q) * PARK_CHUNK + tid; ts[q] = __ldcg(park_ts + g); k[q] = __ldcg(park_key + g); }
        unsigned int pos[SCATTER_WAYS];
#pragma unroll
        for (int q = 0; q < SCATTER_WAYS; ++q) if (on[q]) pos[q] = __ldcg(bounds + (unsigned int)(ts[q] >> 32)) + (unsigned int)ts[q];
#pragma unroll
        for (int q = 0; q < SCATTER_WAYS; ++q) if (on[q]) pairs[pos[q]] = k[q];
    }
}

void launch_scatter_pairs(const FrameCtl* ctl, const unsigned long long* park_key, const unsigned long long* park_ts,
                          const unsigned int* chunk_fill, const unsigned int* bounds, unsigned long long* pairs,
                          void* status_slot, unsigned int frame_no, int sm_count, cudaStream_t s)
{
    launch_pdl(scatter_pairs_kernel, dim3(sm_count * 8), dim3(PARK_CHUNK), 0, s, ctl, park_key, park_ts, chunk_fill, bounds, pairs,
               (volatile FrameStatusSlotDev*)status_slot, frame_no);
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
#pragma unroll
            for (int j = 16; j >= 1; j >>= 1) shuffle_stage<E>(x, tid, j, k);
        }
    }
};
template <int E>
struct UpperLevels<E, 0> { static __device__ __forceinline__ void run(unsigned long long (&)[E], unsigned long long*, unsigned int) {} };

template <int E>
__device__ __forceinline__ void bitonic_sort_regs(unsigned long long (&x)[E], unsigned long long* s_x, unsigned int tid) {
#pragma unroll
    for (int k = 2; k <= 32; k <<= 1)
#pragma unroll
        for (int j = k >> 1; j >= 1; j >>= 1) shuffle_stage<E>(x, tid, j, k);
    exchange_stage<E>(x, s_x, tid, 32, 64);
#pragma unroll
    for (int j = 16; j >= 1; j >>= 1) shuffle_stage<E>(x, tid, j, 64);
    exchange_stage<E>(x, s_x, tid, 64, 128);
    exchange_stage<E>(x, s_x, tid, 32, 128);
#pragma unroll
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
#pragma unroll
            for (int k = 2; k <= 32; k <<= 1)
#pragma unroll
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
