/* libsplat_cuda — C ABI of the B200 splat rasteriser (sm_100a).
 *
 * This is the drop-in boundary for the reference's splat hot path. Where the
 * reference binds Metal buffers/textures and encodes passes
 * (src/instanced_splat_renderer.mm:399-422 upload, :424-528 render,
 * :491-494,507-509 binding table), a host binds these entry points instead.
 * Plain pointers and sizes only; no C++/torch types. All functions return 0 on
 * success or a negative SPLAT_ERR_* code; splat_last_error_string() explains.
 * File:line citations are relative to the reference repository root.
 */
#ifndef SPLAT_CUDA_H
#define SPLAT_CUDA_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SPLAT_TILE 16

enum { SPLAT_OK = 0, SPLAT_ERR_ARG = -1, SPLAT_ERR_CUDA = -2, SPLAT_ERR_OVERFLOW = -3, SPLAT_ERR_STATE = -4 };

/* Renderer mode: which reference shader's per-pixel rule is applied. */
enum {
    SPLAT_MODE_TILE    = 0, /* shaders/gaussian_splat_tile.metal:178-267  (K = 32; 24 per gaussian_splat_types.h:10) */
    SPLAT_MODE_LAYER50 = 1, /* shaders/gaussian_splat_50layer.metal:149-223 (K = 50) */
    SPLAT_MODE_MLAB    = 2  /* shaders/gaussian_splat.metal:201-291 + 329-361: streaming 6-layer K-buffer in
                               instance order, then resolve_main; the framebuffer receives resolve_main's
                               value (src/instanced_splat_renderer.mm:530-577; max_fragments / order unused) */
};
/* Depth order of the per-pixel sort. */
enum {
    SPLAT_ORDER_NEAR_FIRST = 0, /* ascending zFront: what the shader comments intend    */
    SPLAT_ORDER_REFERENCE  = 1  /* descending zFront: what the shader code literally does
                                   (gaussian_splat_tile.metal:244, 50layer.metal:201)  */
};
enum {
    SPLAT_FLAG_TIMING = 1,        /* record per-stage CUDA events (adds event overhead)          */
    SPLAT_FLAG_KEEP_UNSORTED = 2, /* accepted for compatibility, no effect: the unsorted pairs stay fetchable from every frame */
    SPLAT_FLAG_COUNT_WORK = 8,    /* the blend counts the (warp, splat) candidates it executes (splat_status.blend_warp_candidates):
                                     the work count of the blend's roofline; a slower instantiation, not for timed frames */
    SPLAT_FLAG_HOST_BGRA8 = 4     /* splat_render_host / splat_render_batch_host deliver the frame as the
                                     reference's drawable format, BGRA8Unorm (4 B / pixel, packed on the
                                     device by the splat_pack_bgra8 kernel) instead of half4 (8 B / pixel) */
};

/* Mirrors src/instanced_splat_renderer.h:6-11 (56 bytes, no padding). */
typedef struct {
    float rotation[4]; /* w x y z */
    float scale[3];
    float position[3];
    float color[4];    /* r g b opacity */
} SplatInstancePOD;

/* Mirrors shaders/gaussian_splat_types.h:13-17 / gaussian_splat_tile.metal:160-164:
 * the author's 10-byte fragment node (5 halfs). Kept for layout parity; per-pixel
 * lists never leave the chip in this implementation. */
typedef struct { uint16_t depth, color[3], alpha; } SplatFragmentPOD;

typedef struct {
    int32_t mode;           /* SPLAT_MODE_*                                             */
    int32_t max_fragments;  /* K: 1..64 (reference values 24 / 32 / 50); 0 = mode default */
    int32_t order;          /* SPLAT_ORDER_*                                            */
    int32_t sh_degree;      /* 0 = reference parity (src/ply_loader.cpp:9-20); 1..3 extension */
    int32_t tile_row_begin; /* band of tile rows rendered on this device: [begin, end)  */
    int32_t tile_row_end;   /* <= 0: all rows                                           */
    int32_t flags;          /* SPLAT_FLAG_*                                             */
    int32_t reserved;
} splat_opts;

/* 48-byte per-splat record produced by the preprocess kernel (debug fetch). */
typedef struct {
    float cx, cy, ux, uy, vx, vy, opacity;
    uint32_t rg, bd;
    float z_front;
    uint32_t rect_lo, rect_hi;
} splat_record;

typedef struct {
    uint32_t n_splats, n_visible_unused;
    uint64_t n_pairs;       /* D of the last completed frame                            */
    uint64_t pair_capacity;
    uint32_t overflow;      /* 1 if the candidate slots exceeded the capacity: the frame was rendered EMPTY, the buffers
                               have been grown; render it again (n_pairs then holds the candidate count)        */
    uint32_t sort_passes;   /* 1: every tile list is sorted once in shared memory                                */
    uint64_t n_candidates;  /* parking slots handed out for the pairs (whole chunks of 256: >= n_pairs); what the capacity must hold */
    uint32_t max_tile_pairs;/* longest tile list of the last frame                                              */
    uint32_t n_live_blocks; /* 256-splat blocks of the scene that survived block culling (of ceil(n_splats / 256)) */
    uint64_t blend_warp_candidates; /* frames rendered with SPLAT_FLAG_COUNT_WORK: (8x4-pixel warp, splat) tests the blend executed */
} splat_status;

typedef struct splat_ctx splat_ctx;

/* Replaces MTLCreateSystemDefaultDevice + pipeline creation
 * (src/metal_renderer.mm:42-53, src/instanced_splat_renderer.mm:40-222). */
int splat_create(int device, splat_ctx** out);
int splat_destroy(splat_ctx* ctx);

/* Replaces newBufferWithBytes(_instances) (src/instanced_splat_renderer.mm:410-419).
 * aos_host: n SplatInstancePOD records in host memory; converted to SoA planes on the device. */
int splat_upload(splat_ctx* ctx, const void* aos_host, uint32_t n);
/* Extension: f_rest coefficients parsed by src/ply_loader.cpp:76-81,125-127. */
int splat_upload_sh(splat_ctx* ctx, const float* sh_rest45_host, uint32_t n, int degree);

/* Replaces InstancedSplatRenderer::render (src/instanced_splat_renderer.mm:424-528).
 * stream: cudaStream_t (role of the MTLCommandBuffer); fb_half4_dev: device pointer to
 * w*h half4 (role of the drawable). Enqueues and returns without synchronising, exactly as
 * the reference only encodes. view/proj: column-major float[16] (simd_float4x4). */
int splat_render(splat_ctx* ctx, void* stream, void* fb_half4_dev,
                 const float* view16, const float* proj16, int w, int h, const splat_opts* opts);
/* Same, end to end with HOST buffers: uniforms from host, framebuffer copied back to
 * fb_half4_host (pinned or pageable), synchronises, and transparently re-renders if the
 * pair buffers overflowed. */
int splat_render_host(splat_ctx* ctx, void* fb_half4_host,
                      const float* view16, const float* proj16, int w, int h, const splat_opts* opts);
/* Camera-batch: n_views frames, fb_dev[i] device pointers, view/proj 16 floats each. */
int splat_render_batch(splat_ctx* ctx, void* stream, void* const* fb_half4_dev,
                       const float* views16, const float* projs16, int n_views, int w, int h,
                       const splat_opts* opts);

/* Camera batch end to end with HOST buffers: n_views frames rendered back to back on an internal
 * stream while a second stream copies each finished frame to fb_half4_host[i] (pinned memory for
 * real overlap); two device framebuffers alternate. Synchronises before returning and re-runs the
 * batch if the pair buffers overflowed. */
int splat_render_batch_host(splat_ctx* ctx, void* const* fb_half4_host, const float* views16,
                            const float* projs16, int n_views, int w, int h, const splat_opts* opts);

/* After the stream has been synchronised: D, capacity and overflow of the last frame.
 * If overflow is set the buffers have been grown; render the frame again. */
int splat_frame_status(splat_ctx* ctx, splat_status* out);

/* Intermediate stages of the last frame, for bit-exact tests. Synchronises the device.
 * *bytes: in = capacity of host_dst, out = bytes written (or needed if host_dst is NULL). */
enum {
    SPLAT_FETCH_RECORDS = 0,     /* splat_record[n]  (culled entries zeroed)  */
    SPLAT_FETCH_TILES_TOUCHED,   /* uint32[n]                                 */
    SPLAT_FETCH_OFFSETS,         /* uint32[n] exclusive scan                  */
    SPLAT_FETCH_KEYS_UNSORTED,   /* uint64[D]  tile << 32 | depth field, in emission order (instance order, a splat's tiles row-major) */
    SPLAT_FETCH_VALS_UNSORTED,   /* uint32[D]                                 */
    SPLAT_FETCH_KEYS_SORTED,     /* uint64[D]                                 */
    SPLAT_FETCH_VALS_SORTED,     /* uint32[D]                                 */
    SPLAT_FETCH_TILE_RANGES,     /* uint32[T][2]                              */
    SPLAT_FETCH_DIGIT_HISTOGRAMS,/* uint32[8][256]: per-pass digit counts of the last stand-alone sort (frames sort per tile)  */
    SPLAT_FETCH_INSTANCES,       /* SplatInstancePOD[n]: the uploaded scene rebuilt from the device planes (no frame needed) */
    SPLAT_FETCH_SH_REST          /* float[n][45]: uploaded f_rest rows rebuilt from the SH planes (no frame needed)        */
};
int splat_debug_fetch(splat_ctx* ctx, int what, void* host_dst, size_t* bytes);

/* ms of the last frame rendered with SPLAT_FLAG_TIMING:
 * [0] block cull + preprocess + candidate binning [1] tile offsets + scatter [2] per-tile sort [3] unused (was: tile ranges)
 * [4] blend [5] total. Synchronises. */
int splat_stage_times(splat_ctx* ctx, float ms[8]);

/* Stand-alone stable LSD radix sort of (uint64 key, uint32 value) pairs on device
 * buffers (the "sort HBM GB/s" metric). Sorts bits [begin_bit, end_bit) in 8-bit
 * digits; *_alt are scratch of the same size. On return *result_in_alt tells which
 * buffer pair holds the sorted data. Enqueues on stream. */
int splat_sort_pairs_u64_u32(splat_ctx* ctx, void* stream, uint64_t* keys, uint32_t* vals,
                             uint64_t* keys_alt, uint32_t* vals_alt, uint64_t n,
                             int begin_bit, int end_bit, int* result_in_alt);
/* The same sort on 32-bit keys: 16 B per pair and pass instead of 24. end_bit <= 32. */
int splat_sort_pairs_u32_u32(splat_ctx* ctx, void* stream, uint32_t* keys, uint32_t* vals,
                             uint32_t* keys_alt, uint32_t* vals_alt, uint64_t n,
                             int begin_bit, int end_bit, int* result_in_alt);

/* The (uint64 key, uint32 value) sort for callers that cannot allocate: code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc. NO COUNTERPART IN THE REFERENCE.
 * Same kernels and same result as splat_sort_pairs_u64_u32, but it takes no context and calls no cudaMalloc / cudaFree /
 * allocating routine: its scratch (digit histograms, tickets, lookback words) is one fixed array in device memory, about
 * 1 MiB, that every GPU loading this library holds until the process exits. It runs on the CURRENT device (the one `stream`
 * belongs to). n above SPLAT_STATIC_SORT_MAX_PAIRS is refused with SPLAT_ERR_ARG (splat_last_error_string() says why). Every
 * call on a device shares that array (it is cleared on `stream` at the start of each call): calls whose work overlaps on
 * the same GPU - two streams, two threads - race; calls on one stream are safe back to back. */
#define SPLAT_STATIC_SORT_MAX_PAIRS 2097152
int splat_sort_pairs_u64_u32_static(void* stream, uint64_t* keys, uint32_t* vals, uint64_t* keys_alt, uint32_t* vals_alt,
                                    uint64_t n, int begin_bit, int end_bit, int* result_in_alt);

/* GPU-side PLY vertex decode (the step before the path): `rows_host` is the binary body of a PLY vertex
 * element, n_rows rows of stride_floats 4-byte little-endian floats (the reference reads every property as
 * a float, src/ply_loader.cpp:104-110). slot_dst[j] says where the j-th property goes in the reference's
 * PointData (src/ply_loader.h:7-17): 0-2 x y z, 3-5 normals, 6-8 f_dc, 9 opacity, 10-12 scale, 13-16 rot,
 * 17-61 f_rest_0..44, or -1 = ignored (src/ply_loader.cpp:59-81). The kernel applies sigmoid / exp / DC ->
 * colour exactly as src/ply_loader.cpp:112-139 does (bit-identical to the host C library's expf; a NaN stays a
 * NaN but loses its payload), drops rows
 * outside |x|,|y|,|z| < crop_half_extent (src/instanced_splat_renderer.mm:360-366; <= 0 keeps all), keeps
 * file order, and installs the result as the scene (as splat_upload would), with the f_rest coefficients
 * as by splat_upload_sh when sh_degree is 1..3. *n_kept receives the number of splats kept. Synchronous. */
int splat_upload_ply_rows(splat_ctx* ctx, const void* rows_host, uint64_t n_rows, uint32_t stride_floats,
                          const int32_t* slot_dst, float crop_half_extent, int sh_degree, uint32_t* n_kept);

/* Presentation step after the path: RGBA16F framebuffer -> BGRA8Unorm, the format of the reference's
 * drawable (src/metal_renderer.mm:58; blit / resolve at src/instanced_splat_renderer.mm:513-527,557-576).
 * byte = rint(clamp(x, 0, 1) * 255), NaN -> 0; bytes in memory B G R A. Device pointers (16- and 8-byte
 * aligned); enqueues on stream. */
int splat_pack_bgra8(splat_ctx* ctx, void* stream, const void* fb_half4, void* bgra8, int w, int h);

/* Multi-GPU single frame: one context per device, each renders a band of tile rows (splat_opts.tile_row_begin /
 * tile_row_end) into a framebuffer of the full size; the reference has a single Metal device and no counterpart
 * (src/instanced_splat_renderer.mm:424-528).
 * splat_band_rows: contiguous bands [begin, end) of tile rows for `world` ranks into bands[2 * world]; row_weights (one
 * non-negative cost per tile row, e.g. the pairs per row of a previous frame) or NULL for an even split. The boundaries
 * minimise the heaviest band.
 * splat_gather_bands: the path's one exchange step, grouped ncclSend / ncclRecv of the bands' pixel rows
 * (pixel_rows[2 * world] = [begin, end) per rank, w pixels of bytes_per_pixel each) to rank dst's copy of the framebuffer, on
 * `stream`. nccl_comm is the caller's ncclComm_t; NCCL itself is looked up in the calling process at run time. */
int splat_band_rows(int tiles_y, int world, const double* row_weights, int32_t* bands);
int splat_gather_bands(void* nccl_comm, void* stream, void* fb_dev, const int32_t* pixel_rows, int w, int bytes_per_pixel,
                       int rank, int world, int dst);
/* The same exchange for n_fb framebuffers of the same partition (consecutive frames of a pipelined loop) inside ONE NCCL group:
 * one fused launch instead of n_fb. */
int splat_gather_bands_batch(void* nccl_comm, void* stream, void* const* fbs_dev, int n_fb, const int32_t* pixel_rows, int w,
                             int bytes_per_pixel, int rank, int world, int dst);

/* Fused exchange for the banded frame: instead of gathering, every rank's blend kernel writes its band straight into the
 * destination rank's framebuffer over NVLink. The destination allocates its framebuffers with splat_peer_alloc (device memory
 * of the current device + a 64-byte handle to send to the other processes by any means); each other rank maps it with
 * splat_peer_open (current device = its own GPU; peer access is enabled on first use) and passes the mapped pointer to
 * splat_render / splat_render_batch as fb_half4_dev together with its tile_row_begin / tile_row_end. The pixels are in
 * place when the writer's stream has passed the render AND the two GPUs have synchronised (an event the destination
 * waits on, or any collective enqueued after the render on both sides). splat_peer_close unmaps, splat_peer_free frees.
 * No counterpart in the reference (single Metal device). */
int splat_peer_alloc(size_t bytes, void** ptr, void* handle64);
int splat_peer_free(void* ptr);
int splat_peer_open(const void* handle64, void** ptr);
int splat_peer_close(void* ptr);

const char* splat_last_error_string(void);
int splat_version(void);

#ifdef __cplusplus
}
#endif
#endif
