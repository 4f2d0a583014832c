// Internal declarations shared by the kernels and the C ABI (not installed).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include "../../include/splat_cuda.h"

#define SPLAT_MAX_SORT_PASSES 8
#define SORT_THREADS 256
#define SORT_ITEMS 16
#define SORT_ITEMS_SMALL 8
#define SORT_ITEMS_NARROW 24                         // 32-bit keys: 6144 pairs per tile
#define SORT_TILE_KEYS (SORT_THREADS * SORT_ITEMS)   // keys per onesweep tile (large sorts)
#define SORT_TILE_KEYS_SMALL (SORT_THREADS * SORT_ITEMS_SMALL)  // frame-sized sorts

// Per-frame uniforms: role of the reference's Uniforms buffer
// (shaders/gaussian_splat_tile.metal:16-21, src/instanced_splat_renderer.mm:456-469),
// passed as a kernel argument so several frames can be in flight on one stream.
struct FrameUniforms {
    float V[16], P[16], VP[16];
    float W, H;
    int w, h, tiles_x, tiles_y, row_begin, row_end;
    int mode, order, K, sh_degree;
    float cam[3];
    unsigned int n;
    // K1's screening, precomputed on the host (conservative: each carries its own slack): 0.5 W, 0.5 H,
    // (P00 W/2)^2 * B and (P11 H/2)^2 * B with B = 1.02 * a bound on lambda_max(W^T W), W = V's 3x3 block,
    // and the band's pixel rows [scr_ylo, scr_yhi]
    float scr_hw, scr_hh, scr_fx2b, scr_fy2b, scr_ylo, scr_yhi;
};

// Device layout of the records: the 48-byte splat_record padded to one aligned 64-byte row, so K1's scattered record
// writes fill whole 32-byte sectors (a 48-byte stride left partial sectors: read-modify-write in L2 / DRAM, ncu had
// 182 MB of DRAM reads on a band frame whose inputs are 125 MB) and the blend's gathers touch exactly two sectors.
// pad: kx, ky of the D9' tile test (splat_math.cuh: tile_ratios), then two unused words
struct alignas(64) splat_record_row { splat_record r; float kx, ky; unsigned int pad[2]; };

// Device control block of a frame, followed in the same allocation by the per-tile pair counters
// (tile_count[T + 1], then rounded up): one memsetAsync zeroes both at the start of every frame.
struct FrameCtl {
    unsigned long long total_pairs;   // D: pairs that passed the tile test (tile_offsets_kernel)
    unsigned int overflow;            // the parking chunks exceed the pair capacity: the frame's lists are not built
    unsigned int n_chunks;            // parking chunks handed out by K1 (PARK_CHUNK slots each; counted even beyond the capacity)
    unsigned int n_live_blocks;       // 256-splat blocks that survived block culling (block_cull_kernel)
    unsigned int max_tile_pairs;      // longest tile list (tile_offsets_kernel)
    unsigned long long blend_warp_candidates;     // SPLAT_FLAG_COUNT_WORK: (warp, splat) candidates the blend executed
    unsigned int pad[8];
};
static_assert(sizeof(FrameCtl) == 64, "FrameCtl is 64 bytes");
__host__ __device__ __forceinline__ unsigned int* ctl_tile_count(FrameCtl* c) { return reinterpret_cast<unsigned int*>(c + 1); }
#ifndef SCENE_BLOCK
#define SCENE_BLOCK 64u                 // splats per culling block of the Morton-ordered scene (a power of two, 32 .. 256)
#endif
#define PARK_CHUNK 256u               // slots per parking chunk (one warp owns a chunk at a time)

// Control block of the stand-alone radix sort (splat_sort_pairs_*; also sorts the scene's Morton codes at upload).
struct SortCtl {
    unsigned int n_sort;
    unsigned int overflow;            // always 0 (kept so the kernels read one flag)
    unsigned int sort_ticket[SPLAT_MAX_SORT_PASSES];
    unsigned int pad[6];
    unsigned int hist[SPLAT_MAX_SORT_PASSES][256];
};

// SoA planes written once by splat_upload (AoS SplatInstance -> SoA).
struct ScenePlanes {
    const float4* rot;        // w x y z
    const float4* pos_sx;     // position.xyz, scale.x
    const float4* syz_op_rg;  // scale.y, scale.z, opacity, bits(half r | half g << 16)
    const uint16_t* b;        // half(b)
    const float4* base_rgb;   // fp32 r g b (loader colour) for the SH extension; .w unused
    const unsigned int* orig; // original instance index of plane entry i (the planes are in Morton order of the positions)
    const float4* blk_lo;     // per block of 256 entries: min xyz of the finite positions, largest scale
    const float4* blk_hi;     // max xyz
    unsigned int n_blocks;
    const float4* sh;         // SH planes [12][n_pad]: float4 j of splat i holds coefficients f' = 4j..4j+3,
                              // f' = 3*k + channel (coefficient-major, so degree L reads planes 0 .. 3*((L+1)^2-1)/4)
    size_t n_pad;
};

struct SortPlan {
    int n_pass;
    int shift[SPLAT_MAX_SORT_PASSES];   // digit = (key >> shift) & mask
    unsigned int mask[SPLAT_MAX_SORT_PASSES];  // (1 << bits) - 1, bits <= 8
    int slot[SPLAT_MAX_SORT_PASSES];    // which FrameCtl::hist row holds this pass's digit counts
    int ticket[SPLAT_MAX_SORT_PASSES];  // which FrameCtl::sort_ticket this pass draws from
};

// Epoch-tagged 64-bit lookback word: [63:34] epoch, [33:32] flag, [31:0] value.
// Stale words from earlier launches never match the current epoch, so the
// status arrays are never cleared.
#define LB_FLAG_AGG 1ull
#define LB_FLAG_INC 2ull
__device__ __forceinline__ unsigned long long lb_pack(unsigned int epoch, unsigned long long flag, unsigned int v) {
    return ((unsigned long long)epoch << 34) | (flag << 32) | v;
}
__device__ __forceinline__ unsigned long long lb_load(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void lb_store(unsigned long long* p, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// Programmatic dependent launch: the frame's kernels are launched with programmatic stream serialization, so a
// kernel's CTAs become resident and run up to pdl_prologue() while the previous kernel of the stream is still
// draining; griddepcontrol.wait returns when every prerequisite grid has COMPLETED and its writes are visible, so
// the data dependencies are those of a plain in-order stream. Every kernel launched that way calls
// pdl_prologue() first thing, on every thread, before any global access or early return.
__device__ __forceinline__ void pdl_prologue() {
#ifdef __CUDA_ARCH__
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
#endif
}
#ifdef __CUDACC__
#include <cstdlib>
#include <utility>
// Set by the entry point before it enqueues a frame: off for splat_render_batch, whose two frames in flight fill
// each other's gaps already - there the early-resident, waiting CTAs only take slots from the other frame
// (1M / 1080p camera batch 3480 -> 3150 frames/s with it, while a frame on its own gains 3%: 2795 -> 2885).
extern thread_local bool g_pdl_frame;
template <typename... KArgs, typename... Args>
static inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args&&... args) {
    static const bool env_on = [] { const char* e = getenv("SPLAT_PDL"); return !(e && e[0] == '0'); }();   // tuning knob
    const bool on = env_on && g_pdl_frame;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = on ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}
#endif

// Launchers (each enqueues on `s` and returns).
// Scene order at upload: bounding box of the finite positions (as order-preserving unsigned keys: min xyz, max xyz),
// 63-bit Morton codes + identity values for the sort, then the permuted planes + block bounds.
void launch_scene_aabb(const float* aos_dev, unsigned int n, unsigned int* mm6_dev, cudaStream_t s);
void launch_morton(const float* aos_dev, unsigned int n, const float lo[3], const float hi[3], unsigned long long* keys, unsigned int* vals, cudaStream_t s);
void launch_upload_soa(const float* aos_dev, const unsigned int* perm, unsigned int n, float4* rot, float4* pos_sx, float4* syz_op_rg,
                       uint16_t* b, float4* base_rgb, unsigned int* orig, float4* blk_lo, float4* blk_hi, cudaStream_t s);
// sh_rest45_dev: [n][45] as the loader stores it (channel-major: f = channel*15 + k) -> coefficient-major planes
void launch_upload_sh(const float* sh_rest45_dev, const unsigned int* perm, unsigned int n, size_t n_pad, float4* planes, cudaStream_t s);
// K1 (k_preprocess.cu): block culling, then per live block the records (rec[original index]) and the parked candidates.
void launch_preprocess(const ScenePlanes& sp, const FrameUniforms& U, splat_record_row* rec, FrameCtl* ctl, unsigned int* live,
                       unsigned long long* park_key, unsigned long long* park_ts, unsigned int* chunk_fill, unsigned long long capacity,
                       unsigned int live_blocks_hint, cudaStream_t s);      // hint: live blocks of an earlier, similar frame (0 = none): sizes K1's grid
// K2 binning (k_bin.cu). K1 parked every (splat, tile) pair that passed the D9' tile test as (key = depth field << 32 |
// instance index; tile << 32 | slot in the tile's list) and counted the pairs per tile. tile_offsets scans the counters
// into bounds[T + 1]; scatter_pairs moves every parked pair to pairs[bounds[tile] + slot]; tile_sort sorts every tile's
// list in shared memory on the whole 64-bit pair: a total order, so the result equals a stable sort of instance-ordered
// pairs on tile | depth. LAYER50 also gets the list in instance order (arrival).
// order[0 .. band tiles): the band's tile ids, longest lists first: the CTA -> tile mapping of the sort and the blend
void launch_tile_offsets(const FrameUniforms& U, FrameCtl* ctl, unsigned int* bounds, unsigned int* order, cudaStream_t s);
// status_slot: 32 bytes of mapped page-locked memory that receive D / slots used / overflow / longest list and, last, frame_no
void launch_scatter_pairs(const FrameCtl* ctl, const unsigned long long* park_key, const unsigned long long* park_ts,
                          const unsigned int* chunk_fill, const unsigned int* bounds, unsigned long long* pairs,
                          void* status_slot, unsigned int frame_no, int sm_count, cudaStream_t s);
void launch_tile_sort(const FrameUniforms& U, const FrameCtl* ctl, const unsigned int* bounds, const unsigned int* order,
                      unsigned long long* pairs, unsigned int* arrival, cudaStream_t s);
// Stand-alone sort (k_sort.cu). Per-pass digit histograms of keys[0..n) into ctl->hist.
void launch_sort_histogram(const void* keys, int key_bytes, unsigned long long n, SortCtl* ctl, SortPlan plan,
                           int sm_count, cudaStream_t s);
// Sorts n pairs, n read from ctl->n_sort on the device. Result lands in (keys, vals) when
// plan.n_pass is even, in (keys_alt, vals_alt) when odd. Uses epochs epoch_base .. +n_pass-1.
// key_bytes: 4 or 8. n_hint: the host's estimate of n, used only to pick a kernel shape.
void launch_radix_sort(void* keys, unsigned int* vals, void* keys_alt, unsigned int* vals_alt, int key_bytes,
                       unsigned long long n_hint, SortCtl* ctl, unsigned long long* sort_status,
                       unsigned int epoch_base, SortPlan plan, int sm_count, bool small_tiles, cudaStream_t s);
void launch_pack_bgra8(const void* fb_half4, void* bgra8, size_t n_pixels, cudaStream_t s);
cudaError_t launch_ply_decode_chunk(const float* rows_dev, unsigned int n_rows, unsigned int stride, const int* slot_dst_dev,
                                    int jx, int jy, int jz, float crop, unsigned int* block_counts, unsigned int* running,
                                    float* aos_out, float* sh_out, cudaStream_t s);
// pairs: the tile lists (u64 per pair, instance index in the low word), depth-ordered (TILE, LAYER50) or instance-ordered
// (MLAB); arrival: LAYER50's instance-ordered list (u32 per pair). bounds[t] .. bounds[t + 1] = tile t's list.
void launch_blend(const FrameUniforms& U, const splat_record_row* rec, const unsigned long long* pairs,
                  const unsigned int* arrival, const unsigned int* bounds, const unsigned int* order, void* fb_half4, FrameCtl* ctl,
                  bool count_work, cudaStream_t s);
