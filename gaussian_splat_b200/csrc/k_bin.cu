// K2: from K1's parked (splat, tile) pairs to sorted per-tile lists. Replaces the hardware tiler's binning of the
// instanced quads (src/instanced_splat_renderer.mm:496-499 draw call) and the per-pixel insertion sorts of
// shaders/gaussian_splat_tile.metal:240-249 / gaussian_splat_50layer.metal:198-206.
//
//   (K1)                 every (splat, tile of its rectangle) candidate that passes the D9' tile test draws its slot from its
//                        tile's counter and is parked (key = depth field << 32 | instance index; tile, slot) - k_preprocess.cu
//   tile_offsets_kernel  exclusive scan of the T tile counters -> bounds[T + 1], D,
//                        the longest list
//   scatter_pairs_kernel pairs[bounds[tile] + slot] = key
//   tile_sort_kernel     one CTA per tile (longest lists first) sorts its list in the order of the whole 64-bit pair: a
//                        bitonic network on 32-bit keys (half depth << 16 | slot), then runs of equal half depth put in
//                        instance-index order. Instance indices are unique within a list, so that order is total and is
//                        what a stable sort of instance-ordered pairs on tile | depth gives, whatever order
//                        the atomics handed the slots out in.
// Round 1 sorted the frame's (tile << 16 | depth) keys with four onesweep passes over all D pairs and searched the tile
// boundaries afterwards: at frame sizes (1.6M pairs = less than one wave of sort tiles) that was four latency-bound
// chains of load / rank / lookback / scatter, 66 + 10 us of a 346 us frame at 23% of HBM. Here every pair is written
// twice and read twice, and the sort runs on T independent lists of a few hundred pairs.
#include "splat_math.cuh"

#define TILE_CLASSES 24u        // list-length classes of the tile order: 0 = empty, c = lists of 2^(c-1) .. 2^c - 1 pairs

// bounds[t] = first pair of tile t's list, for the tiles [first, first + count] of the frame's band (the blend and the sort
// read nothing else). One CTA: a thread owns a contiguous run of counters, read with up to V independent 16-byte loads (one
// L2 round trip), scanned in registers, written back as vectors. The counter and bounds arrays are padded to
// 65536 + 8 words; the run starts at `first` rounded down to a vector boundary, counters outside the band count as 0.
template <int V>      // 16-byte vectors per thread: bands of up to 4096 V - 4 tiles
__global__ void __launch_bounds__(1024) tile_offsets_kernel(FrameCtl* ctl, unsigned int first, unsigned int count,
                                                            unsigned int* __restrict__ bounds, unsigned int* __restrict__ order)
{
    __shared__ unsigned int s_wsum[32], s_wmax[32], s_cls[TILE_CLASSES];
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
    // order[0 .. count): the band's tiles, longest lists first (classes of list length, one per power of two; the order inside
    // a class is whatever the shared-memory atomics give - nothing depends on it). The sort and the blend take their tiles in
    // this order: in row-major order the long lists of a dense screen region started late and were each kernel's tail
    // (a band of eight of the 6M / 4K frame: sort 34 .. 79 us for the same number of pairs).
    if (tid < TILE_CLASSES) s_cls[tid] = 0;
    __syncthreads();
#pragma unroll
    for (int q = 0; q < V; ++q)
        if ((unsigned)q < per4) {
            const unsigned int* e = &v[q].x;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const unsigned int t = mine + 4u * q + k;
                if (t >= first && t < end) atomicAdd(&s_cls[min(32u - (unsigned int)__clz(e[k]), TILE_CLASSES - 1u)], 1u);
            }
        }
    __syncthreads();
    if (tid == 0) {           // class c starts after every longer class
        unsigned int run2 = 0;
        for (int c = TILE_CLASSES - 1; c >= 0; --c) { const unsigned int x = s_cls[c]; s_cls[c] = run2; run2 += x; }
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < V; ++q)
        if ((unsigned)q < per4) {
            const unsigned int* e = &v[q].x;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const unsigned int t = mine + 4u * q + k;
                if (t >= first && t < end) order[atomicAdd(&s_cls[min(32u - (unsigned int)__clz(e[k]), TILE_CLASSES - 1u)], 1u)] = t;
            }
        }
}

void launch_tile_offsets(const FrameUniforms& U, FrameCtl* ctl, unsigned int* bounds, unsigned int* order, cudaStream_t s)
{
    const unsigned int first = (unsigned int)(U.row_begin * U.tiles_x), count = (unsigned int)(max(U.row_end - U.row_begin, 0) * U.tiles_x);
    const unsigned int span = first + count - (first & ~3u) + 4u;
    if (span <= 16384u) launch_pdl(tile_offsets_kernel<4>, dim3(1), dim3(1024), 0, s, ctl, first, count, bounds, order);
    else if (span <= 32768u) launch_pdl(tile_offsets_kernel<8>, dim3(1), dim3(1024), 0, s, ctl, first, count, bounds, order);
    else launch_pdl(tile_offsets_kernel<16>, dim3(1), dim3(1024), 0, s, ctl, first, count, bounds, order);
}

#define SCATTER_WAYS 8
// A CTA takes eight parking chunks at a time, one slot per thread and chunk: the (key; tile, slot) loads, the bounds lookups
// and the stores are eight independent chains, no atomics (K1 drew the slots). Thread 0 of CTA 0 also leaves the frame's
// status in page-locked host memory (the frame number last), where the next entry-point call finds it without a copy.
struct FrameStatusSlotDev { unsigned long long total_pairs, total_slots; unsigned int overflow, max_tile_pairs, frame_no, live_blocks; };
__global__ void __launch_bounds__(PARK_CHUNK) scatter_pairs_kernel(const FrameCtl* ctl, const unsigned long long* park_key,
                                                                   const unsigned long long* park_ts, const unsigned int* chunk_fill,
                                                                   const unsigned int* bounds, unsigned long long* __restrict__ pairs,
                                                                   volatile FrameStatusSlotDev* slot, unsigned int frame_no)
{
    pdl_prologue();
    const unsigned int over = __ldcg(&ctl->overflow), n_chunks = __ldcg(&ctl->n_chunks);
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        slot->total_pairs = __ldcg(&ctl->total_pairs); slot->total_slots = (unsigned long long)n_chunks * PARK_CHUNK;
        slot->overflow = over; slot->max_tile_pairs = __ldcg(&ctl->max_tile_pairs); slot->live_blocks = __ldcg(&ctl->n_live_blocks);
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
#define TS_MAX_E 16                                   // keys per thread: lists of up to 2048 pairs are sorted in registers
#define TS_CHUNK (TS_THREADS * TS_MAX_E)              // pairs the shared buffers hold
#define TS_MAX_RUN 48u                                // longest run of equal half depths the tie fix-up walks (else: 64-bit network)
#define SENTINEL 0xFFFFFFFFFFFFFFFFull

// Bitonic network over P = 128 E keys of type T, key i = e * 128 + tid held by thread tid in x[e]: partners at distance >= 128
// sit in the same thread (registers), at distance 32 and 64 in another warp (exchanged through shared memory), below that in
// the same warp (shuffles). Position i keeps the smaller key when bit j of i is clear in an ascending block (bit k of i
// clear). Every loop unrolls: register indices, strides and direction bits are compile-time constants.
__device__ __forceinline__ unsigned long long shfl_xor64(unsigned long long v, int m) {
    const unsigned int lo = __shfl_xor_sync(0xFFFFFFFFu, (unsigned int)v, m), hi = __shfl_xor_sync(0xFFFFFFFFu, (unsigned int)(v >> 32), m);
    return ((unsigned long long)hi << 32) | lo;
}
__device__ __forceinline__ unsigned long long shfl_xor_key(unsigned long long v, int m) { return shfl_xor64(v, m); }
__device__ __forceinline__ unsigned int shfl_xor_key(unsigned int v, int m) { return __shfl_xor_sync(0xFFFFFFFFu, v, m); }

// One level (block size K) of the network: its stages j = K/2 .. 1. A template parameter per level keeps every unrolled loop
// small (at most 11 stages x 16 keys); one nested loop over all 66 stages of a 2048-key network was not unrolled and left
// the keys in local memory.
template <typename T, int E, int K>
__device__ __forceinline__ void bitonic_level(T (&x)[E], T* s_x, unsigned int tid) {
    constexpr int W = TS_THREADS;
#pragma unroll
    for (int j = K >> 1; j >= 1; j >>= 1) {
        if (j >= W) {
            // partner e ^ (j / W) in the same thread; bits of i at and above W come from e alone
#pragma unroll
            for (int e = 0; e < E; ++e)
                if ((e & (j / W)) == 0) {
                    const bool asc = ((e * W) & K) == 0;
                    const T a = x[e], b = x[e | (j / W)];
                    const bool sw = (a > b) == asc;
                    x[e] = sw ? b : a; x[e | (j / W)] = sw ? a : b;
                }
        } else if (j >= 32) {
#pragma unroll
            for (int e = 0; e < E; ++e) s_x[e * W + tid] = x[e];
            __syncthreads();
#pragma unroll
            for (int e = 0; e < E; ++e) {
                const unsigned int i = e * W + tid;
                const T p = s_x[i ^ j];
                const bool keep_min = ((i & j) == 0) == ((i & K) == 0);
                x[e] = (keep_min == (x[e] < p)) ? x[e] : p;
            }
            __syncthreads();
        } else {
#pragma unroll
            for (int e = 0; e < E; ++e) {
                const unsigned int i = e * W + tid;
                const T p = shfl_xor_key(x[e], j);
                const bool keep_min = ((i & j) == 0) == ((i & K) == 0);
                x[e] = (keep_min == (x[e] < p)) ? x[e] : p;
            }
        }
    }
}
template <typename T, int E, int K>
struct BitonicLevels {
    static __device__ __forceinline__ void run(T (&x)[E], T* s_x, unsigned int tid) {
        BitonicLevels<T, E, K / 2>::run(x, s_x, tid);
        bitonic_level<T, E, K>(x, s_x, tid);
    }
};
template <typename T, int E>
struct BitonicLevels<T, E, 1> { static __device__ __forceinline__ void run(T (&)[E], T*, unsigned int) {} };

template <typename T, int E>
__device__ __forceinline__ void bitonic_sort_regs(T (&x)[E], T* s_x, unsigned int tid) {
    BitonicLevels<T, E, E * TS_THREADS>::run(x, s_x, tid);
}

// One tile list of 33 .. 128 E pairs. The network runs on 32-bit keys, half the instructions of the 64-bit pairs:
//   depth order   key = half-depth field << 16 | position in the unsorted list. Equal half depths (a few per list: a half has
//                 ~3000 distinct values over a scene's depth range) then sit together in arbitrary order; every element walks
//                 its run of equal depths, counts the members with a smaller instance index and takes that place in the run:
//                 exactly the order of the 64-bit key depth << 32 | instance index. A list with a run longer than
//                 TS_MAX_RUN (a wall of splats at one depth) is sorted once more by the 64-bit network instead.
//   MLAB          pairs carry no depth: the key is the instance index itself.
//   LAYER50       also wants the list in instance (= arrival) order: a second run on the instance indices.
template <int E>
__device__ __forceinline__ void sort_list_fast(unsigned long long* seg, unsigned int* arrival, unsigned int len, bool by_instance,
                                               bool layer, unsigned long long* s_pairs, unsigned int* s_k, unsigned int* s_flag,
                                               unsigned int tid) {
    unsigned int x[E];
    if (tid == 0) *s_flag = 0u;
#pragma unroll
    for (int e = 0; e < E; ++e) {
        const unsigned int i = e * TS_THREADS + tid;
        x[e] = 0xFFFFFFFFu;
        if (i < len) {
            const unsigned long long pr = __ldcg(seg + i);
            s_pairs[i] = pr;
            x[e] = by_instance ? (unsigned int)pr : (((unsigned int)(pr >> 32) << 16) | i);
        }
    }
    __syncthreads();
    bitonic_sort_regs<unsigned int, E>(x, s_k, tid);
    if (by_instance) {
#pragma unroll
        for (int e = 0; e < E; ++e) { const unsigned int i = e * TS_THREADS + tid; if (i < len) seg[i] = (unsigned long long)x[e]; }
        return;
    }
#pragma unroll
    for (int e = 0; e < E; ++e) s_k[e * TS_THREADS + tid] = x[e];
    __syncthreads();
    bool too_long = false;
    for (unsigned int i = tid; i < len; i += TS_THREADS) {          // the thread's own sorted keys, read back from shared memory
        const unsigned int key = s_k[i], d = key >> 16;
        const unsigned long long pr = s_pairs[key & 0xFFFFu];
        const unsigned int idx = (unsigned int)pr;
        unsigned int left = 0, right = 0, rank = 0;
        for (unsigned int j = i; j > 0u && left <= TS_MAX_RUN; --j) {
            const unsigned int kj = s_k[j - 1u];
            if ((kj >> 16) != d) break;
            ++left; rank += (unsigned int)s_pairs[kj & 0xFFFFu] < idx ? 1u : 0u;
        }
        for (unsigned int j = i + 1u; j < len && right <= TS_MAX_RUN; ++j) {
            const unsigned int kj = s_k[j];
            if ((kj >> 16) != d) break;
            ++right; rank += (unsigned int)s_pairs[kj & 0xFFFFu] < idx ? 1u : 0u;
        }
        if (left > TS_MAX_RUN || right > TS_MAX_RUN) too_long = true;
        else seg[i - left + rank] = pr;
    }
    if (too_long) *s_flag = 1u;
    __syncthreads();
    const bool fallback = *s_flag != 0u;
    if (layer) {
#pragma unroll
        for (int e = 0; e < E; ++e) { const unsigned int i = e * TS_THREADS + tid; x[e] = i < len ? (unsigned int)s_pairs[i] : 0xFFFFFFFFu; }
        bitonic_sort_regs<unsigned int, E>(x, s_k, tid);
#pragma unroll
        for (int e = 0; e < E; ++e) { const unsigned int i = e * TS_THREADS + tid; if (i < len) arrival[i] = x[e]; }
    }
    if (fallback) {
        // a long run of equal half depths: the whole pairs through the 64-bit network (s_pairs doubles as its exchange buffer)
        unsigned long long y[E];
#pragma unroll
        for (int e = 0; e < E; ++e) { const unsigned int i = e * TS_THREADS + tid; y[e] = i < len ? s_pairs[i] : SENTINEL; }
        __syncthreads();
        bitonic_sort_regs<unsigned long long, E>(y, s_pairs, tid);
#pragma unroll
        for (int e = 0; e < E; ++e) { const unsigned int i = e * TS_THREADS + tid; if (i < len) seg[i] = y[e]; }
    }
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
                                                               const unsigned int* bounds, const unsigned int* order,
                                                               unsigned long long* pairs, unsigned int* arrival)
{
    __shared__ unsigned long long s_x[TS_CHUNK];           // the list's pairs (and the exchange buffer of the 64-bit network)
    __shared__ unsigned int s_k[TS_CHUNK];                 // exchange buffer of the 32-bit network, then the sorted keys
    __shared__ unsigned int s_flag;
    pdl_prologue();
    const unsigned int tile = __ldcg(order + blockIdx.x);          // longest lists first
    const unsigned int b0 = __ldcg(bounds + tile), len = __ldcg(bounds + tile + 1) - b0;
    const unsigned int tid = threadIdx.x;
    if (len <= 1 && U.mode != SPLAT_MODE_LAYER50) return;
    if (len == 0) return;
    unsigned long long* seg = pairs + b0;
    const bool layer = U.mode == SPLAT_MODE_LAYER50, by_instance = U.mode == SPLAT_MODE_MLAB;
    if (len <= 32) {
        // one warp, shuffles only, on the whole 64-bit pairs
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
    if (len <= 1 * TS_THREADS) sort_list_fast<1>(seg, arrival + b0, len, by_instance, layer, s_x, s_k, &s_flag, tid);
    else if (len <= 2 * TS_THREADS) sort_list_fast<2>(seg, arrival + b0, len, by_instance, layer, s_x, s_k, &s_flag, tid);
    else if (len <= 4 * TS_THREADS) sort_list_fast<4>(seg, arrival + b0, len, by_instance, layer, s_x, s_k, &s_flag, tid);
    else if (len <= 8 * TS_THREADS) sort_list_fast<8>(seg, arrival + b0, len, by_instance, layer, s_x, s_k, &s_flag, tid);
    else if (len <= 16 * TS_THREADS) sort_list_fast<16>(seg, arrival + b0, len, by_instance, layer, s_x, s_k, &s_flag, tid);
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

void launch_tile_sort(const FrameUniforms& U, const FrameCtl* ctl, const unsigned int* bounds, const unsigned int* order,
                      unsigned long long* pairs, unsigned int* arrival, cudaStream_t s)
{
    const int rows = U.row_end - U.row_begin;
    if (rows <= 0 || U.tiles_x <= 0) return;
    launch_pdl(tile_sort_kernel, dim3((unsigned int)(U.tiles_x * rows)), dim3(TS_THREADS), 0, s, U, ctl, bounds, order, pairs, arrival);
}
