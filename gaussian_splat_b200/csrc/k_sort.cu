// Stable LSD "onesweep" radix sort of (key, uint32 value) pairs, 8-bit digits: the stand-alone sort entry points
// (splat_sort_pairs_u64_u32 / _u32_u32: 24 / 16 B per pair and pass) and the Morton ordering of the scene at upload.
// Frames no longer sort their pairs globally (k_bin.cu: per-tile lists sorted in shared memory). One kernel per digit pass; each pass reads and writes every pair exactly once. Per-warp
// digit ranks come from eight ballots per key, tile prefixes from an epoch-tagged decoupled lookback per digit. Replaces the per-pixel insertion sorts of
// shaders/gaussian_splat_tile.metal:240-249 / gaussian_splat_50layer.metal:198-206 with
// one global (tile | depth) sort.
#include "splat_internal.cuh"
#include <cstdlib>

#define SORT_WARPS (SORT_THREADS / 32)
#ifndef LB_DEPTH
#define LB_DEPTH 4
#endif
// -DSORT_PROFILE: thread 0 of every CTA accumulates clock64() per phase into ctl->hist[7][0..7] (tuning aid).
#ifdef SORT_PROFILE
#define PROF(k) do { if (tid == 0) { const long long t_ = clock64(); prof_acc[k] += t_ - prof_last; prof_last = t_; } } while (0)
#else
#define PROF(k) do {} while (0)
#endif

template <typename KeyT>
__global__ void __launch_bounds__(256) sort_histogram_kernel(const KeyT* __restrict__ keys,
                                                             unsigned long long n, SortCtl* ctl,
                                                             const __grid_constant__ SortPlan plan)
{
    __shared__ unsigned int h[SPLAT_MAX_SORT_PASSES][256];
    for (int p = 0; p < SPLAT_MAX_SORT_PASSES; ++p) h[p][threadIdx.x] = 0;
    __syncthreads();
    // four independent key loads in flight per thread (one at a time left the pass at 1.7 TB/s: 0.29 ms of a 4.05 ms sort of
    // 64M pairs), then their 4 x n_pass shared-memory atomics, which bound it now (8 and 16 loads in flight: no faster)
    const unsigned long long stride = (unsigned long long)gridDim.x * 256ull;
    unsigned long long i = (unsigned long long)blockIdx.x * 256ull + threadIdx.x;
    for (; i + 3ull * stride < n; i += 4ull * stride) {
        KeyT k[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) k[u] = __ldg(keys + i + (unsigned long long)u * stride);
        for (int p = 0; p < plan.n_pass; ++p) {
            const int sh = plan.shift[p]; const unsigned int m = plan.mask[p]; unsigned int* hp = h[plan.slot[p]];
#pragma unroll
            for (int u = 0; u < 4; ++u) atomicAdd(hp + ((unsigned int)(k[u] >> sh) & m), 1u);
        }
    }
    for (; i < n; i += stride) {
        const KeyT k = __ldg(keys + i);
        for (int p = 0; p < plan.n_pass; ++p) atomicAdd(&h[plan.slot[p]][(unsigned int)(k >> plan.shift[p]) & plan.mask[p]], 1u);
    }
    __syncthreads();
    for (int p = 0; p < SPLAT_MAX_SORT_PASSES; ++p) {
        const unsigned int v = h[p][threadIdx.x];
        if (v) atomicAdd(&ctl->hist[p][threadIdx.x], v);
    }
}

void launch_sort_histogram(const void* keys, int key_bytes, unsigned long long n, SortCtl* ctl, SortPlan plan,
                           int sm_count, cudaStream_t s)
{
    if (n == 0) return;
    unsigned long long blocks = (n + 256ull * 16 - 1) / (256ull * 16);
    if (blocks > (unsigned long long)sm_count * 8) blocks = (unsigned long long)sm_count * 8;
    if (key_bytes == 4) sort_histogram_kernel<unsigned int><<<(unsigned int)blocks, 256, 0, s>>>((const unsigned int*)keys, n, ctl, plan);
    else sort_histogram_kernel<unsigned long long><<<(unsigned int)blocks, 256, 0, s>>>((const unsigned long long*)keys, n, ctl, plan);
}

__device__ __forceinline__ void bulk_prefetch_l2(const void* gptr, unsigned int bytes) {   // bytes % 16 == 0
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(gptr), "r"(bytes) : "memory");
}

// peers &= lanes whose digit agrees with mine in the bit `bitmask`: test, vote, select, one three-input logic op
// (ptxas sets the eight predicates of a digit with one R2P).
__device__ __forceinline__ unsigned int ballot_match_step(unsigned int peers, unsigned int d, unsigned int bitmask) {
    asm("{\n\t.reg .pred q;\n\t.reg .b32 t, m, nb;\n\t"
        "and.b32 t, %1, %2;\n\tsetp.ne.u32 q, t, 0;\n\t"
        "vote.sync.ballot.b32 m, q, 0xffffffff;\n\t"
        "selp.b32 nb, 0, 0xffffffff, q;\n\t"
        "lop3.b32 %0, %0, m, nb, 0x60;\n\t}"          // peers & (m ^ nb)
        : "+r"(peers) : "r"(d), "r"(bitmask));
    return peers;
}

// Exclusive scan of one value per thread over a 256-thread block. s_w: SORT_WARPS words.
__device__ __forceinline__ unsigned int block_excl_scan_256(unsigned int v, unsigned int* s_w, unsigned int& total) {
    const unsigned int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned int incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { unsigned int t = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= d) incl += t; }
    __syncthreads();            // protect s_w reuse
    if (lane == 31) s_w[warp] = incl;
    __syncthreads();
    unsigned int base = 0; total = 0;
#pragma unroll
    for (int k = 0; k < SORT_WARPS; ++k) { const unsigned int t = s_w[k]; if (k < (int)warp) base += t; total += t; }
    return base + incl - v;
}

struct SortShared { unsigned int gexcl[256], off[256], w[SORT_WARPS], tile[2]; };

template <int ITEMS, typename KeyT, int RANK_SETS>
__device__ __forceinline__ void onesweep_pass_body(
    const KeyT* __restrict__ keys_in, const unsigned int* __restrict__ vals_in,
    KeyT* __restrict__ keys_out, unsigned int* __restrict__ vals_out,
    SortCtl* ctl, unsigned long long* status, unsigned int epoch, int pass, int tix, int shift, unsigned int mask,
    unsigned long long* s_raw, SortShared& sh)
{
    constexpr int TILE_KEYS = SORT_THREADS * ITEMS;
    KeyT* s_keys = reinterpret_cast<KeyT*>(s_raw);                 // [TILE_KEYS] keys, then values
    unsigned int* const s_gexcl = sh.gexcl; unsigned int* const s_off = sh.off; unsigned int* const s_w = sh.w;
    unsigned int* const s_tile = sh.tile;
    unsigned int* s_vals = reinterpret_cast<unsigned int*>(s_keys + TILE_KEYS);
    // per set, warp, digit: keys of that digit seen so far in the warp (later: tile-local offsets)
    unsigned short (*s_wh)[SORT_WARPS][256] = reinterpret_cast<unsigned short (*)[SORT_WARPS][256]>(s_vals + TILE_KEYS);

    if (ctl->overflow) return;
    const unsigned int n = ctl->n_sort;
    const unsigned int n_tiles = (n + TILE_KEYS - 1) / TILE_KEYS;
    const unsigned int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned int lt_mask = (1u << lane) - 1u;

    {   // bucket bases from the global digit histogram of this pass
        unsigned int tot;
        s_gexcl[tid] = block_excl_scan_256(ctl->hist[pass][tid], s_w, tot);
    }

#ifdef SORT_PROFILE
    long long prof_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0}, prof_last = clock64();
#endif
    // Tickets are drawn one tile ahead so the atomic's round trip hides behind the current tile.
    // A CTA works through its tickets in increasing order, so every predecessor of a tile it waits
    // on in the lookback is owned by a running CTA: no deadlock.
    unsigned int next_ticket = 0;
    if (tid == 0) next_ticket = atomicAdd(&ctl->sort_ticket[tix], 1u);
    unsigned int parity = 0;
    while (true) {
        // one barrier per tile: the ticket slot is double-buffered, so thread 0 may publish the next tile's
        // number while slower warps still read the current one (it is rewritten two barriers later)
        if (tid == 0) s_tile[parity] = next_ticket;
        PROF(6);
        __syncthreads();
        const unsigned int tile = s_tile[parity];
        parity ^= 1u;
        PROF(0);
        if (tile >= n_tiles) break;
        const unsigned int tbase = tile * TILE_KEYS;
        const unsigned int tcount = min((unsigned int)TILE_KEYS, n - tbase);

        // ---- load (warp-striped: order (warp, item, lane) == memory order)
        KeyT key[ITEMS];
        unsigned int val[ITEMS];
        const unsigned int wofs = warp * (32 * ITEMS) + lane;
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) {
            const unsigned int li = wofs + i * 32;
            key[i] = li < tcount ? __ldcg(keys_in + tbase + li) : ~(KeyT)0;      // L2 only: the buffer was written by the previous pass (PDL-linked)
        }
        if (tid == 0) {
            // next ticket, and its tile pulled into L2 by the TMA unit while this tile is processed: the
            // plain loads above otherwise wait ~10K cycles on DRAM with nothing else of this CTA to run
            next_ticket = atomicAdd(&ctl->sort_ticket[tix], 1u);
            if (next_ticket < n_tiles) {
                const unsigned int nb = next_ticket * TILE_KEYS;
                const unsigned int nc = min((unsigned int)TILE_KEYS, n - nb) & ~3u;
                if (nc) {
                    bulk_prefetch_l2(keys_in + nb, nc * (unsigned int)sizeof(KeyT));
                    bulk_prefetch_l2(vals_in + nb, nc * 4u);
                }
            }
        }
#pragma unroll
        for (int k = 0; k < 4; ++k)     // 256 16-bit counters per (set, warp) = 128 words
#pragma unroll
            for (int c = 0; c < RANK_SETS; ++c) reinterpret_cast<unsigned int*>(&s_wh[c][warp][0])[lane + k * 32] = 0u;
        __syncwarp();
#ifdef SORT_PROFILE
        { unsigned long long x_ = 0;
#pragma unroll
          for (int i = 0; i < ITEMS; ++i) x_ ^= key[i] ^ val[i];
          if (x_ == 0xdeadbeefcafe1234ull) prof_acc[7]++; }
        PROF(1);
#endif

        // ---- per-warp ranks. Peer mask of a key = AND over the digit's eight bits of (ballot of that bit,
        // complemented where the key's bit is clear): registers and the vote unit only. The running count of
        // the digit in this warp lives in a 16-bit shared counter: every lane reads it, the lowest peer adds
        // the round's peers. History at 64M pairs: hardware match.any saturated the ADU pipe (70%, about 64
        // cycles per MATCH); a shared-memory bitmap match (atomicOr, load, clearing store per key, 4 counter
        // sets) held the LSU pipe 81% busy, an ATOMS costing about two LSU cycles per lane; the ballots need no
        // bitmap, so the LSU pipe drops to 61% and one counter set (least clearing) is the quickest:
        // 2.92 -> 3.19 TB/s, and the frame's four passes 77 -> 73 us.
        // RANK_SETS independent counter sets (items [c*ITEMS/SETS, (c+1)*ITEMS/SETS) use set c) share the two
        // warp barriers of a round.
        // ranks are below TILE_KEYS <= 8192: two per register (every index below is a compile-time constant)
        unsigned int rankp[(ITEMS + 1) / 2];
#pragma unroll
        for (int i = 0; i < (ITEMS + 1) / 2; ++i) rankp[i] = 0u;
#define RANK_GET(i) ((rankp[(i) >> 1] >> (((i) & 1) * 16)) & 0xFFFFu)
#define RANK_SET(i, v) (rankp[(i) >> 1] = (rankp[(i) >> 1] & ~(0xFFFFu << (((i) & 1) * 16))) | ((v) << (((i) & 1) * 16)))
        constexpr int PER_SET = ITEMS / RANK_SETS;
#pragma unroll
        for (int i = 0; i < PER_SET; ++i) {
            unsigned int dd[RANK_SETS], peers[RANK_SETS], before[RANK_SETS];
#pragma unroll
            for (int c = 0; c < RANK_SETS; ++c) {
                dd[c] = (unsigned int)(key[c * PER_SET + i] >> shift) & mask;
                before[c] = *reinterpret_cast<volatile unsigned short*>(&s_wh[c][warp][dd[c]]);
                unsigned int p = 0xFFFFFFFFu;
#pragma unroll
                for (int b = 0; b < 8; ++b) p = ballot_match_step(p, dd[c], 1u << b);
                peers[c] = p;
                RANK_SET(c * PER_SET + i, before[c] + __popc(p & lt_mask));
            }
            __syncwarp();
#pragma unroll
            for (int c = 0; c < RANK_SETS; ++c)
                if ((peers[c] & lt_mask) == 0)
                    *reinterpret_cast<volatile unsigned short*>(&s_wh[c][warp][dd[c]]) = (unsigned short)(before[c] + __popc(peers[c]));
            __syncwarp();
        }
        __syncthreads();
        PROF(2);

        // ---- digit column: totals over warps, tile-local exclusive scan, lookback
        unsigned int cnt = 0;
#pragma unroll
        for (int w = 0; w < SORT_WARPS; ++w)
#pragma unroll
            for (int c = 0; c < RANK_SETS; ++c) cnt += s_wh[c][w][tid];
        const unsigned int valid_cnt = cnt - (tid == mask ? (TILE_KEYS - tcount) : 0u);
        unsigned long long* st = status + (size_t)tile * 256 + tid;
        lb_store(st, lb_pack(epoch, tile == 0 ? LB_FLAG_INC : LB_FLAG_AGG, valid_cnt));
        unsigned int tot;
        const unsigned int lexcl = block_excl_scan_256(cnt, s_w, tot);
        {
            unsigned int run = lexcl;
#pragma unroll
            for (int w = 0; w < SORT_WARPS; ++w)
#pragma unroll
                for (int c = 0; c < RANK_SETS; ++c) {
                    const unsigned int t = s_wh[c][w][tid]; s_wh[c][w][tid] = (unsigned short)run; run += t;
                }
        }
        __syncthreads();
        PROF(3);
        // ---- pairs: registers -> shared at their tile-local rank (frees the registers before the
        // lookback; global offsets are not needed yet)
        // keys first: once they sit in shared memory their registers are free, and the values - only needed now - are loaded
        // into them (from L2: the tile was prefetched there one tile ahead). Values held through the ranking cost a register
        // per pair; without them, and with the ranks packed two per register, a 64-bit tile grows from 16 to 24 pairs per
        // thread at the same two CTAs per SM, which amortises the per-tile chain (counter clearing, digit scans, on average
        // 3.6 serial L2 round trips of lookback): 64M pairs 3.31 -> 3.92 TB/s. Measured around it: 20 / 22 / 26 / 28 pairs per
        // thread 3.63 / 3.88 / 3.81 / 3.58 (spills from 26 on); lookback depth 3 / 6 / 8 at 24 pairs 3.91 / 3.88 / 3.79.
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) {
            const unsigned int d = (unsigned int)(key[i] >> shift) & mask;
            const unsigned int r = RANK_GET(i) + (unsigned int)s_wh[i / PER_SET][warp][d];
            RANK_SET(i, r);
            s_keys[r] = key[i];
        }
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) {
            const unsigned int li = wofs + i * 32;
            val[i] = li < tcount ? __ldcg(vals_in + tbase + li) : 0u;
        }
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) s_vals[RANK_GET(i)] = val[i];
        PROF(4);
        // Decoupled lookback, LB_DEPTH predecessors in flight per step: with every tile of a small
        // sort resident at once, a one-at-a-time walk is a serial chain of L2 round trips.
        unsigned int prefix = 0;
        for (int look = (int)tile - 1; look >= 0; look -= LB_DEPTH) {
            unsigned long long wd[LB_DEPTH];
#pragma unroll
            for (int j = 0; j < LB_DEPTH; ++j)
                if (look - j >= 0) wd[j] = lb_load(status + (size_t)(look - j) * 256 + tid);
            bool stop = false;
#pragma unroll
            for (int j = 0; j < LB_DEPTH; ++j) {
                if (look - j >= 0 && !stop) {
                    while ((unsigned int)(wd[j] >> 34) != epoch || ((wd[j] >> 32) & 3ull) == 0)
                        wd[j] = lb_load(status + (size_t)(look - j) * 256 + tid);
                    prefix += (unsigned int)wd[j];
                    stop = ((wd[j] >> 32) & 3ull) == LB_FLAG_INC;
                }
            }
            if (stop) break;
        }
        if (tile != 0) lb_store(st, lb_pack(epoch, LB_FLAG_INC, prefix + valid_cnt));
        s_off[tid] = s_gexcl[tid] + prefix - lexcl;
        __syncthreads();
        PROF(5);

        // ---- shared (tile-sorted) -> global (contiguous per digit run)
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) {
            const unsigned int j = tid + i * SORT_THREADS;
            if (j < tcount) {
                const KeyT k = s_keys[j];
                const unsigned int g = s_off[(unsigned int)(k >> shift) & mask] + j;
                keys_out[g] = k;
                vals_out[g] = s_vals[j];
            }
        }
    }
#ifdef SORT_PROFILE
    if (tid == 0) for (int k = 0; k < 8; ++k) atomicAdd(&ctl->hist[7][k], (unsigned int)(prof_acc[k] >> 4));
#endif
}

// SETS / SETS_LONG rank-counter sets: with different values, sorts of more than SORT_LONG pairs would take the
// SETS_LONG instantiation on the device's own count. Measured 6% slower on long sorts than a kernel with the one
// body (more spills), so the callers below pass SETS == SETS_LONG and choose on the host's estimate of n.
#define SORT_LONG 8000000u
#ifndef K_ITEMS_WIDE
#define K_ITEMS_WIDE 24           // 64-bit keys: 6144 pairs per tile (see the staging comment in the pass body)
#endif
#ifndef K_CTAS_WIDE
#define K_CTAS_WIDE 2
#endif
#ifndef K_CTAS_NARROW
#define K_CTAS_NARROW 2
#endif
#ifndef K_ITEMS_NARROW
#define K_ITEMS_NARROW SORT_ITEMS_NARROW
#endif
#ifndef SORT_SETS
#define SORT_SETS 1          // 32-bit keys, frame-sized sorts
#endif
#ifndef SORT_SETS_LONG
#define SORT_SETS_LONG 1     // 32-bit keys, more than SORT_LONG pairs
#endif
#ifndef SORT_SETS_WIDE
#define SORT_SETS_WIDE 1     // 64-bit keys
#endif
template <int ITEMS, typename KeyT, int SETS, int SETS_LONG, int MIN_CTAS>
__global__ void __launch_bounds__(SORT_THREADS, MIN_CTAS) onesweep_pass_kernel(
    const KeyT* __restrict__ keys_in, const unsigned int* __restrict__ vals_in,
    KeyT* __restrict__ keys_out, unsigned int* __restrict__ vals_out,
    SortCtl* ctl, unsigned long long* status, unsigned int epoch, int pass, int tix, int shift, unsigned int mask)
{
    extern __shared__ unsigned long long s_raw[];
    __shared__ SortShared sh;
    pdl_prologue();
    if (SETS_LONG != SETS && ctl->n_sort > SORT_LONG)
        onesweep_pass_body<ITEMS, KeyT, SETS_LONG>(keys_in, vals_in, keys_out, vals_out, ctl, status, epoch, pass, tix, shift, mask, s_raw, sh);
    else
        onesweep_pass_body<ITEMS, KeyT, SETS>(keys_in, vals_in, keys_out, vals_out, ctl, status, epoch, pass, tix, shift, mask, s_raw, sh);
}

template <int ITEMS, typename KeyT, int SETS, int SETS_LONG, int MIN_CTAS>
static void run_passes(KeyT* ki, unsigned int* vi, KeyT* ko, unsigned int* vo, SortCtl* ctl,
                       unsigned long long* sort_status, unsigned int epoch_base, const SortPlan& plan, int grid,
                       cudaStream_t s)
{
    constexpr int smem = SORT_THREADS * ITEMS * ((int)sizeof(KeyT) + 4) + (SETS > SETS_LONG ? SETS : SETS_LONG) * SORT_WARPS * 256 * 2;
    // function attributes are per device: set on every call (cheap), so a second context on another device works
    cudaFuncSetAttribute(onesweep_pass_kernel<ITEMS, KeyT, SETS, SETS_LONG, MIN_CTAS>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    for (int p = 0; p < plan.n_pass; ++p) {
        launch_pdl(onesweep_pass_kernel<ITEMS, KeyT, SETS, SETS_LONG, MIN_CTAS>, dim3(grid), dim3(SORT_THREADS), smem, s,
                   (const KeyT*)ki, (const unsigned int*)vi, ko, vo, ctl, sort_status, epoch_base + p,
                   plan.slot[p], plan.ticket[p], plan.shift[p], plan.mask[p]);
        KeyT* tk = ki; ki = ko; ko = tk;
        unsigned int* tv = vi; vi = vo; vo = tv;
    }
}

// small_tiles: 2048-pair tiles, 3 CTAs per SM. Measured slower than 4096-pair tiles at every size tried
// (1.6M, 12M, 64M pairs: the per-tile fixed costs dominate), so callers pass false; kept as a tuning
// knob (SPLAT_SORT_TILE=small|large).
void launch_radix_sort(void* keys, unsigned int* vals, void* keys_alt, unsigned int* vals_alt, int key_bytes,
                       unsigned long long n_hint, SortCtl* ctl, unsigned long long* sort_status,
                       unsigned int epoch_base, SortPlan plan, int sm_count, bool small_tiles, cudaStream_t s)
{
    if (key_bytes == 4) {
        // 6144-pair tiles (24 per thread; measured best of 2048 .. 8192: the per-tile fixed costs - counter
        // clearing, digit scans, lookback - amortise, beyond that registers spill).
        unsigned int* k0 = (unsigned int*)keys; unsigned int* k1 = (unsigned int*)keys_alt;
        if (n_hint > SORT_LONG) run_passes<K_ITEMS_NARROW, unsigned int, SORT_SETS_LONG, SORT_SETS_LONG, K_CTAS_NARROW>(k0, vals, k1, vals_alt, ctl, sort_status, epoch_base, plan, sm_count * K_CTAS_NARROW, s);
        else run_passes<K_ITEMS_NARROW, unsigned int, SORT_SETS, SORT_SETS, K_CTAS_NARROW>(k0, vals, k1, vals_alt, ctl, sort_status, epoch_base, plan, sm_count * K_CTAS_NARROW, s);
        return;
    }
    static const char* force = getenv("SPLAT_SORT_TILE");    // tuning knob: "small" | "large"
    if (force && force[0] == 's') small_tiles = true;
    if (force && force[0] == 'l') small_tiles = false;
    unsigned long long* k0 = (unsigned long long*)keys; unsigned long long* k1 = (unsigned long long*)keys_alt;
    if (small_tiles) run_passes<SORT_ITEMS_SMALL, unsigned long long, 4, 4, 3>(k0, vals, k1, vals_alt, ctl, sort_status, epoch_base, plan, sm_count * 3, s);
    else run_passes<K_ITEMS_WIDE, unsigned long long, SORT_SETS_WIDE, SORT_SETS_WIDE, K_CTAS_WIDE>(k0, vals, k1, vals_alt, ctl, sort_status, epoch_base, plan, sm_count * K_CTAS_WIDE, s);
}
