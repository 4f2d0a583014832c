// libsplat_cuda C ABI (include/splat_cuda.h). Host side of the hot path: owns all device
// scratch, builds the per-frame uniforms (src/instanced_splat_renderer.mm:453-469) and
// enqueues K1..K5 on the caller's stream without synchronising, as the reference's
// render() only encodes (src/instanced_splat_renderer.mm:424-528).
#include "splat_internal.cuh"
#include <nvtx3/nvToolsExt.h>      // header-only NVTX 3: ranges cost nothing unless a tool is attached

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

static thread_local std::string g_err;
// for entry points that live in other files (sort_static.cu): sets this thread's splat_last_error_string()
void splat_set_last_error(const char* msg) { g_err = msg ? msg : ""; }

#define CK(call)                                                                          \
    do {                                                                                  \
        cudaError_t e_ = (call);                                                          \
        if (e_ != cudaSuccess) {                                                          \
            g_err = std::string(#call) + ": " + cudaGetErrorString(e_);                   \
            return SPLAT_ERR_CUDA;                                                        \
        }                                                                                 \
    } while (0)
#define ARG_FAIL(msg) do { g_err = (msg); return SPLAT_ERR_ARG; } while (0)

__global__ void set_n_sort_kernel(SortCtl* ctl, unsigned int n) { ctl->n_sort = n; }

// Everything one frame in flight writes: two such sets let the camera-batch entry points keep two frames
// in flight on two streams (at frame sizes the binning stages are latency-bound and leave most SMs idle). The single
// frame entry points use set 0 only; the second set is allocated on the first batch call.
#define CTL_BYTES (sizeof(FrameCtl) + (65536 + 8) * 4)      // control block + tile counters of the largest frame (65536 tiles)
struct FrameSet {
    splat_record_row* rec = nullptr;        // [n_pad] records, indexed by ORIGINAL instance index
    unsigned int* live = nullptr;           // [n_blocks] blocks that survived block culling
    // parked pairs (K1): park_key / park_ts[cap] in chunks of PARK_CHUNK slots, chunk_fill[cap / PARK_CHUNK];
    // tile lists: pairs[cap] (depth field << 32 | instance index, sorted per tile) and LAYER50's instance-ordered copy
    // arrival[cap_arr] (allocated on the first LAYER50 frame)
    unsigned long long *pairs = nullptr, *park_key = nullptr, *park_ts = nullptr; unsigned int* chunk_fill = nullptr;
    unsigned long long capacity = 0;
    uint32_t* arrival = nullptr; unsigned long long capacity_arr = 0;
    FrameCtl* ctl = nullptr;                // + tile counters (CTL_BYTES)
    uint32_t* bounds = nullptr;             // [65536 + 8]: first pair of every tile's list
    uint32_t* order = nullptr;              // [65536 + 8]: the band's tiles, longest lists first
    cudaStream_t own_stream = nullptr;
};

// Status of frame `frame_no`, written by its tile_offsets_kernel straight into page-locked host memory (no copy is
// enqueued): the next entry-point call sees an overflow without synchronising.
struct FrameStatusSlot { unsigned long long total_pairs, total_slots; unsigned int overflow, max_tile_pairs, frame_no, live_blocks; };
#define STATUS_RING 256

#define FB_RING 4
struct splat_ctx : FrameSet {
    FrameSet alt; bool alt_ready = false; int cur_set = 0;
    int device = 0, sm_count = 148;
    uint32_t n = 0; size_t n_pad = 0;
    float4 *rot = nullptr, *pos_sx = nullptr, *syz = nullptr, *base_rgb = nullptr, *sh = nullptr; uint16_t* bq = nullptr;
    unsigned int* orig = nullptr; float4 *blk_lo = nullptr, *blk_hi = nullptr; uint32_t n_blocks = 0;     // Morton order + block bounds
    int sh_degree = 0;
    FrameCtl* ctl_host = nullptr;
    FrameStatusSlot* status_ring = nullptr;      // page-locked, mapped
    uint32_t frame_no = 0, checked_frame = 0;    // frames enqueued / frames whose status has been looked at
    // live blocks of the latest landed frame with the same band and viewport (sizes K1's grid; 0 = no hint)
    uint32_t hint_live = 0; int hint_w = 0, hint_h = 0, hint_r0 = 0, hint_r1 = 0;
    struct { int w, h, r0, r1; } frame_sig[STATUS_RING] = {};
    // stand-alone sort
    SortCtl* sort_ctl = nullptr; unsigned long long* sort_status = nullptr; size_t sort_status_tiles = 0; uint32_t sort_epoch = 0;
    FrameUniforms lastU{}; bool have_frame = false;
    cudaEvent_t ev[6] = {}; bool timed = false;
    void* fb_scratch = nullptr; size_t fb_scratch_bytes = 0;
    // splat_render_batch_host: a ring of FB_RING device framebuffers, so rendering runs
    // ahead of the device-to-host copies instead of waiting for the copy two frames back
    cudaStream_t copy_stream = nullptr; void* fb_ring[FB_RING] = {}; size_t fb_ring_bytes = 0;
    void* bgra_scratch[FB_RING] = {}; size_t bgra_bytes = 0;      // SPLAT_FLAG_HOST_BGRA8 staging
    cudaEvent_t ev_rendered[FB_RING] = {}, ev_copied[FB_RING] = {}, ev_fork = nullptr;
    unsigned int* batch_flags = nullptr; int batch_flags_cap = 0;   // splat_render_batch_host: the frame number of every view
};

static void swap_sets(splat_ctx* c) {
    FrameSet tmp = *static_cast<FrameSet*>(c);
    *static_cast<FrameSet*>(c) = c->alt;
    c->alt = tmp;
    c->cur_set ^= 1;
}
static void use_set(splat_ctx* c, int k) { if (c->cur_set != k) swap_sets(c); }

static void mat4_mul_host(const float* a, const float* b, float* o) {
    // simd_mul(a, b), k = 0..3 left to right (column-major)
    float t[16];
    for (int j = 0; j < 4; ++j)
        for (int r = 0; r < 4; ++r) {
            volatile float acc = a[0 * 4 + r] * b[j * 4 + 0];
            volatile float m1 = a[1 * 4 + r] * b[j * 4 + 1]; acc = acc + m1;
            volatile float m2 = a[2 * 4 + r] * b[j * 4 + 2]; acc = acc + m2;
            volatile float m3 = a[3 * 4 + r] * b[j * 4 + 3]; acc = acc + m3;
            t[j * 4 + r] = acc;
        }
    std::memcpy(o, t, 64);
}

static int free_pairs(splat_ctx* c) {
    cudaFree(c->pairs); cudaFree(c->park_key); cudaFree(c->park_ts); cudaFree(c->chunk_fill); cudaFree(c->arrival);
    c->pairs = c->park_key = c->park_ts = nullptr; c->chunk_fill = nullptr; c->arrival = nullptr;
    c->capacity = 0; c->capacity_arr = 0;
    return 0;
}

static int ensure_sort_status(splat_ctx* c, unsigned long long n_keys) {
    const size_t tiles = (size_t)((n_keys + SORT_TILE_KEYS_SMALL - 1) / SORT_TILE_KEYS_SMALL) + 1;
    if (!c->sort_ctl) { CK(cudaMalloc(&c->sort_ctl, sizeof(SortCtl))); CK(cudaMemset(c->sort_ctl, 0, sizeof(SortCtl))); }
    if (tiles <= c->sort_status_tiles) return 0;
    CK(cudaDeviceSynchronize());
    cudaFree(c->sort_status); c->sort_status = nullptr;
    CK(cudaMalloc(&c->sort_status, tiles * 256 * 8));
    CK(cudaMemset(c->sort_status, 0, tiles * 256 * 8));
    c->sort_status_tiles = tiles;
    return 0;
}

// Pair capacity = parking slots (whole chunks of PARK_CHUNK; a warp's last chunk is partly used, so the pairs are fewer).
// Nothing needs clearing: every slot a kernel reads was written earlier in the same frame.
static int ensure_pairs(splat_ctx* c, unsigned long long cap) {
    if (cap <= c->capacity) return 0;
    if (cap >= 0xFFFF0000ull) ARG_FAIL("pair capacity would exceed 2^32");
    CK(cudaDeviceSynchronize());
    free_pairs(c);
    cap = (cap + 4095) / 4096 * 4096;
    CK(cudaMalloc(&c->pairs, cap * 8)); CK(cudaMalloc(&c->park_key, cap * 8)); CK(cudaMalloc(&c->park_ts, cap * 8));
    CK(cudaMalloc(&c->chunk_fill, cap / PARK_CHUNK * 4));
    c->capacity = cap;
    return 0;
}
static int alloc_frame_fixed(splat_ctx* c) {        // control block + tile counters and bounds of the CURRENT set
    CK(cudaMalloc(&c->ctl, CTL_BYTES)); CK(cudaMemset(c->ctl, 0, CTL_BYTES));
    CK(cudaMalloc(&c->bounds, (65536 + 8) * 4)); CK(cudaMemset(c->bounds, 0, (65536 + 8) * 4));
    CK(cudaMalloc(&c->order, (65536 + 8) * 4)); CK(cudaMemset(c->order, 0, (65536 + 8) * 4));
    return 0;
}

extern "C" {

const char* splat_last_error_string(void) { return g_err.c_str(); }
int splat_version(void) { return 100; }

int splat_create(int device, splat_ctx** out) {
    if (!out) ARG_FAIL("splat_create: out is NULL");
    *out = nullptr;
    int count = 0;
    CK(cudaGetDeviceCount(&count));
    if (device < 0 || device >= count) ARG_FAIL("splat_create: no such CUDA device");
    CK(cudaSetDevice(device));
    splat_ctx* c = new splat_ctx();
    c->device = device;
    CK(cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, device));
    { const int rc = alloc_frame_fixed(c); if (rc) return rc; }
    CK(cudaMallocHost(&c->ctl_host, sizeof(FrameCtl)));
    std::memset(c->ctl_host, 0, sizeof(FrameCtl));
    CK(cudaHostAlloc(&c->status_ring, STATUS_RING * sizeof(FrameStatusSlot), cudaHostAllocMapped | cudaHostAllocPortable));
    std::memset(c->status_ring, 0, STATUS_RING * sizeof(FrameStatusSlot));
    for (int k = 0; k < 6; ++k) CK(cudaEventCreate(&c->ev[k]));
    CK(cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
    CK(cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
    for (int k = 0; k < FB_RING; ++k) {
        CK(cudaEventCreateWithFlags(&c->ev_rendered[k], cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&c->ev_copied[k], cudaEventDisableTiming));
    }
    *out = c;
    return SPLAT_OK;
}

// Per-splat scratch of the CURRENT frame set (records, live-block list).
static void free_scene_scratch(splat_ctx* c) {
    cudaFree(c->rec); cudaFree(c->live);
    c->rec = nullptr; c->live = nullptr;
}
static int alloc_scene_scratch(splat_ctx* c, size_t n_pad) {
    CK(cudaMalloc(&c->rec, n_pad * sizeof(splat_record_row)));
    CK(cudaMemset(c->rec, 0, n_pad * sizeof(splat_record_row)));
    CK(cudaMalloc(&c->live, (n_pad / SCENE_BLOCK + 4) * 4));
    return SPLAT_OK;
}
// Releases the second frame set completely (it is rebuilt by the next camera-batch call).
static void drop_alt(splat_ctx* c) {
    use_set(c, 0);
    if (!c->alt_ready) return;
    swap_sets(c);
    free_scene_scratch(c); free_pairs(c);
    cudaFree(c->bounds); cudaFree(c->order); cudaFree(c->ctl);
    if (c->own_stream) cudaStreamDestroy(c->own_stream);
    *static_cast<FrameSet*>(c) = FrameSet{};
    swap_sets(c);
    c->alt_ready = false;
}
static int ensure_alt(splat_ctx* c) {
    use_set(c, 0);
    if (c->alt_ready) return SPLAT_OK;
    const unsigned long long cap = c->capacity;
    swap_sets(c);
    int rc = SPLAT_OK;
    do {
        if (alloc_frame_fixed(c) != 0 ||
            cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking) != cudaSuccess) { g_err = "second frame set: allocation failed"; rc = SPLAT_ERR_CUDA; break; }
        if (c->n_pad) { rc = alloc_scene_scratch(c, c->n_pad); if (rc) break; }
        rc = ensure_pairs(c, cap ? cap : (1ull << 20));
    } while (0);
    swap_sets(c);
    if (rc == SPLAT_OK) c->alt_ready = true;
    return rc;
}

// Grows the pair buffers of every live frame set; leaves the current set selected.
static int ensure_pairs_all(splat_ctx* c, unsigned long long cap) {
    int rc = ensure_pairs(c, cap);
    if (rc == SPLAT_OK && c->alt_ready) { swap_sets(c); rc = ensure_pairs(c, cap); swap_sets(c); }
    return rc;
}

static void free_scene(splat_ctx* c) {
    drop_alt(c);
    cudaFree(c->rot); cudaFree(c->pos_sx); cudaFree(c->syz); cudaFree(c->bq); cudaFree(c->sh); cudaFree(c->base_rgb);
    cudaFree(c->orig); cudaFree(c->blk_lo); cudaFree(c->blk_hi);
    free_scene_scratch(c);
    c->rot = c->pos_sx = c->syz = c->base_rgb = c->sh = c->blk_lo = c->blk_hi = nullptr; c->bq = nullptr; c->orig = nullptr; c->n_blocks = 0;
    c->n = 0; c->n_pad = 0; c->sh_degree = 0;
}

int splat_destroy(splat_ctx* c) {
    if (!c) return SPLAT_OK;
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    free_scene(c); free_pairs(c);
    cudaFree(c->bounds); cudaFree(c->order); cudaFree(c->ctl); cudaFree(c->fb_scratch); cudaFree(c->sort_ctl); cudaFree(c->sort_status);
    cudaFreeHost(c->ctl_host); cudaFreeHost(c->status_ring);
    for (int k = 0; k < 6; ++k) if (c->ev[k]) cudaEventDestroy(c->ev[k]);
    if (c->own_stream) cudaStreamDestroy(c->own_stream);
    if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
    for (int k = 0; k < FB_RING; ++k) { if (c->ev_rendered[k]) cudaEventDestroy(c->ev_rendered[k]); if (c->ev_copied[k]) cudaEventDestroy(c->ev_copied[k]); }
    if (c->ev_fork) cudaEventDestroy(c->ev_fork);
    for (int k = 0; k < FB_RING; ++k) { cudaFree(c->fb_ring[k]); cudaFree(c->bgra_scratch[k]); }
    std::free(c->batch_flags);
    delete c;
    return SPLAT_OK;
}

// Installs n SplatInstance rows that already sit in device memory as the scene (planes + per-frame scratch).
static int sort_pairs_impl(splat_ctx* c, void* stream, void* keys, uint32_t* vals, void* keys_alt, uint32_t* vals_alt,
                           uint64_t n, int begin_bit, int end_bit, int key_bytes, int* result_in_alt);
static float ordered_to_float(unsigned int k) { k ^= (k >> 31) ? 0x80000000u : 0xFFFFFFFFu; float f; std::memcpy(&f, &k, 4); return f; }

// Installs n SplatInstance rows that already sit in device memory as the scene: Morton order of the positions (our own
// radix sort on 63-bit codes), planes written in that order with the original index beside them, block bounds,
// per-frame scratch. SPLAT_SCENE_ORDER=0 keeps instance order (tuning / debugging aid; results are the same).
static int install_scene(splat_ctx* c, const float* aos_dev, uint32_t n) {
    const size_t n_pad = ((size_t)n + 1023) / 1024 * 1024;
    CK(cudaMalloc(&c->rot, n_pad * 16)); CK(cudaMalloc(&c->pos_sx, n_pad * 16)); CK(cudaMalloc(&c->syz, n_pad * 16));
    CK(cudaMalloc(&c->bq, n_pad * 2)); CK(cudaMalloc(&c->base_rgb, n_pad * 16)); CK(cudaMalloc(&c->orig, n_pad * 4));
    c->n_blocks = (uint32_t)((n + SCENE_BLOCK - 1u) / SCENE_BLOCK);
    CK(cudaMalloc(&c->blk_lo, ((size_t)c->n_blocks + 1) * 16)); CK(cudaMalloc(&c->blk_hi, ((size_t)c->n_blocks + 1) * 16));
    { const int rc = alloc_scene_scratch(c, n_pad); if (rc) return rc; }
    static const bool ordered = [] { const char* e = getenv("SPLAT_SCENE_ORDER"); return !(e && e[0] == '0'); }();
    unsigned long long *keys = nullptr, *keys_alt = nullptr; unsigned int *vals = nullptr, *vals_alt = nullptr, *mm = nullptr;
    const unsigned int* perm = nullptr;
    int rc = SPLAT_OK;
    if (ordered && n > 256) {
        do {
            if (cudaMalloc(&keys, (size_t)n * 8) != cudaSuccess || cudaMalloc(&keys_alt, (size_t)n * 8) != cudaSuccess ||
                cudaMalloc(&vals, (size_t)n * 4) != cudaSuccess || cudaMalloc(&vals_alt, (size_t)n * 4) != cudaSuccess ||
                cudaMalloc(&mm, 32) != cudaSuccess) { g_err = "scene order: allocation failed"; rc = SPLAT_ERR_CUDA; break; }
            launch_scene_aabb(aos_dev, n, mm, 0);
            unsigned int h[6];
            if (cudaMemcpy(h, mm, 24, cudaMemcpyDeviceToHost) != cudaSuccess) { g_err = "scene order: bounding box failed"; rc = SPLAT_ERR_CUDA; break; }
            float lo[3], hi[3];
            for (int a = 0; a < 3; ++a) { lo[a] = ordered_to_float(h[a]); hi[a] = ordered_to_float(h[3 + a]); if (!(lo[a] <= hi[a])) { lo[a] = 0.0f; hi[a] = 0.0f; } }
            launch_morton(aos_dev, n, lo, hi, keys, vals, 0);
            int alt = 0;
            rc = sort_pairs_impl(c, nullptr, keys, vals, keys_alt, vals_alt, n, 0, 64, 8, &alt);
            if (rc) break;
            perm = alt ? vals_alt : vals;
        } while (0);
    }
    if (rc == SPLAT_OK) {
        launch_upload_soa(aos_dev, perm, n, c->rot, c->pos_sx, c->syz, c->bq, c->base_rgb, c->orig, c->blk_lo, c->blk_hi, 0);
        if (cudaGetLastError() != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) { g_err = "scene upload kernels failed"; rc = SPLAT_ERR_CUDA; }
    }
    cudaFree(keys); cudaFree(keys_alt); cudaFree(vals); cudaFree(vals_alt); cudaFree(mm);
    if (rc) return rc;
    c->n = n; c->n_pad = n_pad;
    unsigned long long want = (unsigned long long)n * 12ull;
    if (want < (1ull << 20)) want = 1ull << 20;
    if (want > 0xC0000000ull) want = 0xC0000000ull;
    return ensure_pairs_all(c, want);
}

static int install_sh(splat_ctx* c, const float* sh_dev, uint32_t n, int degree) {
    CK(cudaMalloc(&c->sh, c->n_pad * 12 * 16));
    launch_upload_sh(sh_dev, c->orig, n, c->n_pad, c->sh, 0);
    CK(cudaGetLastError());
    CK(cudaDeviceSynchronize());
    c->sh_degree = degree;
    return SPLAT_OK;
}

int splat_upload(splat_ctx* c, const void* aos_host, uint32_t n) {
    if (!c) ARG_FAIL("splat_upload: ctx is NULL");
    if (n && !aos_host) ARG_FAIL("splat_upload: aos_host is NULL");
    CK(cudaSetDevice(c->device));
    CK(cudaDeviceSynchronize());
    free_scene(c);
    c->have_frame = false;
    if (n == 0) return SPLAT_OK;
    float* aos_dev = nullptr;
    CK(cudaMalloc(&aos_dev, (size_t)n * 56));
    int rc = cudaMemcpy(aos_dev, aos_host, (size_t)n * 56, cudaMemcpyHostToDevice) == cudaSuccess ? SPLAT_OK : SPLAT_ERR_CUDA;
    if (rc) g_err = "splat_upload: host to device copy failed";
    else rc = install_scene(c, aos_dev, n);
    cudaFree(aos_dev);
    return rc;
}

int splat_upload_ply_rows(splat_ctx* c, const void* rows_host, uint64_t n_rows, uint32_t stride, const int32_t* slot_dst,
                          float crop, int sh_degree, uint32_t* n_kept) {
    if (!c) ARG_FAIL("splat_upload_ply_rows: ctx is NULL");
    if (n_kept) *n_kept = 0;
    if (n_rows && (!rows_host || !slot_dst)) ARG_FAIL("splat_upload_ply_rows: NULL rows or slot table");
    if (n_rows >= 0xFFFF0000ull) ARG_FAIL("splat_upload_ply_rows: too many rows");
    if (n_rows && (stride == 0 || stride > 512)) ARG_FAIL("splat_upload_ply_rows: stride must be 1..512 floats");
    if (sh_degree < 0 || sh_degree > 3) ARG_FAIL("splat_upload_ply_rows: sh_degree must be 0..3");
    int jx = -1, jy = -1, jz = -1;                       // a later property of the same name overrides an earlier one
    for (uint32_t j = 0; j < stride && n_rows; ++j) {
        if (slot_dst[j] < -1 || slot_dst[j] > 61) ARG_FAIL("splat_upload_ply_rows: slot index outside -1..61");
        if (slot_dst[j] == 0) jx = (int)j; else if (slot_dst[j] == 1) jy = (int)j; else if (slot_dst[j] == 2) jz = (int)j;
    }
    CK(cudaSetDevice(c->device));
    CK(cudaDeviceSynchronize());
    free_scene(c);
    c->have_frame = false;
    if (n_rows == 0) return SPLAT_OK;

    const uint32_t chunk_rows = 1u << 18;                // 65 MB of a 62-float file per chunk
    const uint32_t max_blocks = (chunk_rows + 127) / 128;
    float *rows_dev = nullptr, *aos_dev = nullptr, *sh_dev = nullptr;
    int* slots_dev = nullptr; unsigned int *counts = nullptr, *running = nullptr;
    int rc = SPLAT_OK;
    auto fail = [&](const char* what) { g_err = std::string("splat_upload_ply_rows: ") + what; rc = SPLAT_ERR_CUDA; };
    const size_t chunk_bytes = (size_t)(n_rows < chunk_rows ? n_rows : chunk_rows) * stride * 4;
    if (cudaMalloc(&rows_dev, chunk_bytes) != cudaSuccess || cudaMalloc(&aos_dev, (size_t)n_rows * 56) != cudaSuccess ||
        (sh_degree > 0 && cudaMalloc(&sh_dev, (size_t)n_rows * 45 * 4) != cudaSuccess) ||
        cudaMalloc(&slots_dev, stride * 4) != cudaSuccess || cudaMalloc(&counts, max_blocks * 4) != cudaSuccess ||
        cudaMalloc(&running, 4) != cudaSuccess) fail("device allocation failed");
    if (!rc && (cudaMemcpy(slots_dev, slot_dst, stride * 4, cudaMemcpyHostToDevice) != cudaSuccess ||
                cudaMemset(running, 0, 4) != cudaSuccess)) fail("slot table copy failed");
    for (uint64_t r0 = 0; !rc && r0 < n_rows; r0 += chunk_rows) {
        const uint32_t nr = (uint32_t)(n_rows - r0 < chunk_rows ? n_rows - r0 : chunk_rows);
        if (cudaMemcpyAsync(rows_dev, (const char*)rows_host + r0 * stride * 4, (size_t)nr * stride * 4, cudaMemcpyHostToDevice,
                            c->own_stream) != cudaSuccess) { fail("host to device copy failed"); break; }
        if (launch_ply_decode_chunk(rows_dev, nr, stride, slots_dev, jx, jy, jz, crop, counts, running, aos_dev, sh_dev,
                                    c->own_stream) != cudaSuccess) { fail(cudaGetErrorString(cudaGetLastError())); break; }
    }
    unsigned int kept = 0;
    if (!rc && (cudaStreamSynchronize(c->own_stream) != cudaSuccess ||
                cudaMemcpy(&kept, running, 4, cudaMemcpyDeviceToHost) != cudaSuccess)) fail(cudaGetErrorString(cudaGetLastError()));
    if (!rc && kept) {
        rc = install_scene(c, aos_dev, kept);
        if (!rc && sh_degree > 0) rc = install_sh(c, sh_dev, kept, sh_degree);
    }
    cudaFree(rows_dev); cudaFree(aos_dev); cudaFree(sh_dev); cudaFree(slots_dev); cudaFree(counts); cudaFree(running);
    if (!rc && n_kept) *n_kept = kept;
    return rc;
}

int splat_upload_sh(splat_ctx* c, const float* sh_host, uint32_t n, int degree) {
    if (!c) ARG_FAIL("splat_upload_sh: ctx is NULL");
    if (degree < 0 || degree > 3) ARG_FAIL("splat_upload_sh: degree must be 0..3");
    if (n != c->n) ARG_FAIL("splat_upload_sh: n differs from splat_upload");
    CK(cudaSetDevice(c->device));
    CK(cudaDeviceSynchronize());
    cudaFree(c->sh); c->sh = nullptr; c->sh_degree = 0;
    if (degree == 0 || n == 0) return SPLAT_OK;
    if (!sh_host) ARG_FAIL("splat_upload_sh: sh_rest45_host is NULL");
    float* staging = nullptr;
    CK(cudaMalloc(&staging, (size_t)n * 45 * 4));
    int rc = cudaMemcpy(staging, sh_host, (size_t)n * 45 * 4, cudaMemcpyHostToDevice) == cudaSuccess ? SPLAT_OK : SPLAT_ERR_CUDA;
    if (rc) g_err = "splat_upload_sh: host to device copy failed";
    else rc = install_sh(c, staging, n, degree);
    cudaFree(staging);
    return rc;
}

}  // extern "C"

// --------------------------------------------------------------------- frames
static int build_frame(splat_ctx* c, const float* view, const float* proj, int w, int h, const splat_opts* o,
                       FrameUniforms& U) {
    if (!view || !proj) ARG_FAIL("splat_render: view/proj is NULL");
    if (w <= 0 || h <= 0 || w > 65536 || h > 65536) ARG_FAIL("splat_render: bad viewport");
    splat_opts d{}; if (o) d = *o;
    if (d.mode != SPLAT_MODE_TILE && d.mode != SPLAT_MODE_LAYER50 && d.mode != SPLAT_MODE_MLAB) ARG_FAIL("splat_render: bad mode");
    if (d.order != SPLAT_ORDER_NEAR_FIRST && d.order != SPLAT_ORDER_REFERENCE) ARG_FAIL("splat_render: bad order");
    int K = d.max_fragments;
    if (K == 0) K = d.mode == SPLAT_MODE_LAYER50 ? 50 : 32;
    if (K < 1 || K > 64) ARG_FAIL("splat_render: max_fragments must be 1..64");
    if (d.mode == SPLAT_MODE_MLAB) K = 0x7FFFFFFF;      // the K-buffer never closes a pixel: every arrival is merged
    if (d.sh_degree < 0 || d.sh_degree > 3) ARG_FAIL("splat_render: sh_degree must be 0..3");
    if (d.sh_degree > c->sh_degree) ARG_FAIL("splat_render: sh_degree exceeds uploaded SH data");
    std::memcpy(U.V, view, 64); std::memcpy(U.P, proj, 64);
    mat4_mul_host(proj, view, U.VP);                    // src/instanced_splat_renderer.mm:453
    U.w = w; U.h = h; U.W = (float)w; U.H = (float)h;
    U.tiles_x = (w + SPLAT_TILE - 1) / SPLAT_TILE; U.tiles_y = (h + SPLAT_TILE - 1) / SPLAT_TILE;
    U.row_begin = 0; U.row_end = U.tiles_y;
    if (d.tile_row_end > 0) {
        if (d.tile_row_begin < 0 || d.tile_row_begin > d.tile_row_end) ARG_FAIL("splat_render: bad tile row band");
        U.row_begin = d.tile_row_begin; U.row_end = d.tile_row_end < U.tiles_y ? d.tile_row_end : U.tiles_y;
        if (U.row_begin > U.row_end) U.row_begin = U.row_end;
    }
    U.mode = d.mode; U.order = d.order; U.K = K; U.sh_degree = d.sh_degree;
    for (int j = 0; j < 3; ++j) {                      // camera position = -R^T t of the view matrix (SH view direction)
        volatile float m0 = view[j * 4 + 0] * view[12], m1 = view[j * 4 + 1] * view[13], m2 = view[j * 4 + 2] * view[14];
        volatile float acc = m0 + m1; acc = acc + m2;
        U.cam[j] = -acc;
    }
    U.n = c->n;
    {   // K1's conservative screening bound: lambda_max(W^T W) <= the largest absolute row sum of W^T W
        // (Gershgorin), W = the view matrix's 3x3 block; 1 for a rigid view. 2% slack for the rounding here.
        float g = 0.0f;
        for (int a = 0; a < 3; ++a) {
            float rs = 0.0f;
            for (int b = 0; b < 3; ++b) {
                float d = 0.0f;
                for (int r = 0; r < 3; ++r) d += view[a * 4 + r] * view[b * 4 + r];
                rs += fabsf(d);
            }
            g = fmaxf(g, rs);
        }
        const float B = g * 1.02f, Fx = proj[0] * (U.W * 0.5f), Fy = proj[5] * (U.H * 0.5f);
        U.scr_hw = 0.5f * U.W; U.scr_hh = 0.5f * U.H;
        U.scr_fx2b = Fx * Fx * B * 1.0001f; U.scr_fy2b = Fy * Fy * B * 1.0001f;
        U.scr_ylo = (float)(U.row_begin * 16);
        U.scr_yhi = fminf((float)(U.row_end * 16), U.H);
    }
    if ((long long)U.tiles_x * U.tiles_y > 65536) ARG_FAIL("splat_render: more than 65536 tiles");
    if (U.tiles_x > 4096 || U.tiles_y > 4096) ARG_FAIL("splat_render: viewport wider or taller than 65536 pixels");
    return SPLAT_OK;
}

// Looks at the status slots of the frames enqueued since the last look (their tile_offsets_kernel writes them into
// page-locked memory): an overflowed frame has built no lists (its framebuffer is empty), so the buffers are grown now
// and the caller is told once, SPLAT_ERR_OVERFLOW, to render again. Frames still in flight are looked at next time.
// wait = true: the device has been synchronised, every enqueued frame has landed.
static int check_overflow(splat_ctx* c, bool wait, unsigned long long* worst_out = nullptr) {
    unsigned long long worst = 0;
    if (c->frame_no - c->checked_frame > STATUS_RING) c->checked_frame = c->frame_no - STATUS_RING;     // older slots are overwritten
    while (c->checked_frame != c->frame_no) {
        const uint32_t f = c->checked_frame + 1;
        const volatile FrameStatusSlot* sl = c->status_ring + (f % STATUS_RING);
        if (sl->frame_no != f) { if (!wait) break; c->checked_frame = f; continue; }    // not landed (or a frame without kernels)
        if (sl->overflow && sl->total_slots > worst) worst = sl->total_slots;
        if (!sl->overflow) {           // a landed frame: its live-block count sizes K1's grid for the next frame of the same shape
            const auto& sg = c->frame_sig[f % STATUS_RING];
            c->hint_live = sl->live_blocks; c->hint_w = sg.w; c->hint_h = sg.h; c->hint_r0 = sg.r0; c->hint_r1 = sg.r1;
        }
        c->checked_frame = f;
    }
    if (worst_out) *worst_out = worst;
    if (!worst) return SPLAT_OK;
    const int set = c->cur_set;
    use_set(c, 0);
    int rc = ensure_pairs(c, worst + worst / 4 + 4096);
    if (rc == SPLAT_OK && c->alt_ready) { swap_sets(c); rc = ensure_pairs(c, worst + worst / 4 + 4096); swap_sets(c); }
    use_set(c, set);
    if (rc) return rc;
    g_err = "an earlier frame overflowed the pair buffers and was rendered empty; the buffers have been grown: render it again";
    return SPLAT_ERR_OVERFLOW;
}

thread_local bool g_pdl_frame = true;
// Internal flag (not a splat_opts flag): launch this frame's kernels without programmatic dependent launch.
#define FRAME_NO_PDL (1 << 30)
static int enqueue_frame(splat_ctx* c, cudaStream_t s, void* fb, const FrameUniforms& U, int flags) {
    g_pdl_frame = (flags & FRAME_NO_PDL) == 0;
    const unsigned int T = (unsigned int)(U.tiles_x * U.tiles_y);
    const bool timed = (flags & SPLAT_FLAG_TIMING) != 0;
    const bool layer = U.mode == SPLAT_MODE_LAYER50;
    if (layer && c->capacity_arr < c->capacity) {
        CK(cudaDeviceSynchronize());
        cudaFree(c->arrival); c->arrival = nullptr;
        CK(cudaMalloc(&c->arrival, c->capacity * 4));
        c->capacity_arr = c->capacity;
    }
    c->frame_no = c->frame_no + 1;
    if (c->frame_no == 0) { c->frame_no = 1; c->checked_frame = 0; }
    FrameStatusSlot* slot = c->status_ring + (c->frame_no % STATUS_RING);
    c->frame_sig[c->frame_no % STATUS_RING] = {U.w, U.h, U.row_begin, U.row_end};
    const uint32_t live_hint = (c->hint_w == U.w && c->hint_h == U.h && c->hint_r0 == U.row_begin && c->hint_r1 == U.row_end) ? c->hint_live : 0u;
    if (timed) CK(cudaEventRecord(c->ev[0], s));
    CK(cudaMemsetAsync(c->ctl, 0, sizeof(FrameCtl) + ((size_t)T + 1) * 4, s));       // control block + tile counters
    // stage labels for Nsight (the reference labels its encoders: src/instanced_splat_renderer.mm:487,504)
    nvtxRangePushA("splat: K1 block cull + preprocess + candidates");
    if (c->n) {
        ScenePlanes sp{c->rot, c->pos_sx, c->syz, c->bq, c->base_rgb, c->orig, c->blk_lo, c->blk_hi, c->n_blocks, c->sh, c->n_pad};
        launch_preprocess(sp, U, c->rec, c->ctl, c->live, c->park_key, c->park_ts, c->chunk_fill, c->capacity, live_hint, s);
    }
    nvtxRangePop();
    if (timed) CK(cudaEventRecord(c->ev[1], s));
    nvtxRangePushA("splat: K2 tile offsets + scatter");
    launch_tile_offsets(U, c->ctl, c->bounds, c->order, s);
    if (c->n) launch_scatter_pairs(c->ctl, c->park_key, c->park_ts, c->chunk_fill, c->bounds, c->pairs, slot, c->frame_no, c->sm_count, s);
    else { slot->total_pairs = 0; slot->total_slots = 0; slot->overflow = 0; slot->max_tile_pairs = 0; slot->frame_no = c->frame_no; }
    nvtxRangePop();
    if (timed) CK(cudaEventRecord(c->ev[2], s));
    nvtxRangePushA("splat: K3 per-tile sort");
    if (c->n) launch_tile_sort(U, c->ctl, c->bounds, c->order, c->pairs, c->arrival, s);
    nvtxRangePop();
    if (timed) { CK(cudaEventRecord(c->ev[3], s)); CK(cudaEventRecord(c->ev[4], s)); }
    nvtxRangePushA("splat: K5 blend");
    launch_blend(U, c->rec, c->pairs, c->arrival, c->bounds, c->order, fb, c->ctl, (flags & SPLAT_FLAG_COUNT_WORK) != 0, s);
    nvtxRangePop();
    if (timed) CK(cudaEventRecord(c->ev[5], s));
    CK(cudaGetLastError());
    c->lastU = U; c->have_frame = true; c->timed = timed;
    return SPLAT_OK;
}

extern "C" {

int splat_render(splat_ctx* c, void* stream, void* fb, const float* view, const float* proj, int w, int h,
                 const splat_opts* o) {
    if (!c) ARG_FAIL("splat_render: ctx is NULL");
    if (!fb) ARG_FAIL("splat_render: framebuffer is NULL");
    CK(cudaSetDevice(c->device));
    use_set(c, 0);
    FrameUniforms U;
    int rc = build_frame(c, view, proj, w, h, o, U);
    if (rc) return rc;
    if (c->capacity == 0) { rc = ensure_pairs(c, 1ull << 20); if (rc) return rc; }
    rc = check_overflow(c, false);          // an earlier frame that overflowed: buffers grown, reported once
    if (rc) return rc;
    return enqueue_frame(c, (cudaStream_t)stream, fb, U, o ? o->flags : 0);
}

// Two frames in flight: the latency-bound stages of one frame (block culling, the one-CTA tile scan, the sort's tail) leave
// SMs idle that the other frame's kernels fill. Measured on B200 with this round's pipeline (device-resident camera batch,
// frames/s serial -> two in flight): 1M / 1080p TILE 3456 -> 3834, 2M LAYER50 1395 -> 1542, 6M / 4K 639 -> 689, one band of
// eight of the 6M / 4K frame 238 -> 197 us per frame. Round 1's pipeline (persistent onesweep CTAs, 16 MB+ working sets
// fighting over shared memory and L2) lost 14 .. 48% on the large cases and kept them serial; that no longer holds.
// SPLAT_TWO_FRAMES=0 / 1 forces either way (tuning knob).
static bool two_frames_in_flight(const splat_ctx*, int, int, const splat_opts*) {
    static const int force = [] { const char* e = getenv("SPLAT_TWO_FRAMES"); return e ? (e[0] == '0' ? 0 : 1) : -1; }();
    return force != 0;
}

// Camera batches keep two frames in flight: view v runs on frame set v & 1 (own scratch, own stream), so the
// latency-bound stages of one frame (K2-K4) fill the SMs the other frame's tails leave idle.
int splat_render_batch(splat_ctx* c, void* stream, void* const* fbs, const float* views, const float* projs,
                       int n_views, int w, int h, const splat_opts* o) {
    if (!c) ARG_FAIL("splat_render_batch: ctx is NULL");
    if (n_views < 0 || (n_views && (!fbs || !views || !projs))) ARG_FAIL("splat_render_batch: bad arguments");
    if (n_views == 0) return SPLAT_OK;
    CK(cudaSetDevice(c->device));
    if (n_views == 1 || !two_frames_in_flight(c, w, h, o)) {
        for (int v = 0; v < n_views; ++v) {
            const int rc = splat_render(c, stream, fbs[v], views + 16 * v, projs + 16 * v, w, h, o);
            if (rc) return rc;
        }
        return SPLAT_OK;
    }
    use_set(c, 0);
    if (c->capacity == 0) { const int rc = ensure_pairs(c, 1ull << 20); if (rc) return rc; }
    { const int rc = ensure_alt(c); if (rc) return rc; }
    { const int rc = check_overflow(c, false); if (rc) return rc; }
    const int flags = o ? (o->flags & ~(SPLAT_FLAG_TIMING | SPLAT_FLAG_KEEP_UNSORTED)) : 0;
    cudaStream_t user = (cudaStream_t)stream;
    CK(cudaEventRecord(c->ev_fork, user));
    for (int v = 0; v < n_views; ++v) {
        if (!fbs[v]) ARG_FAIL("splat_render_batch: NULL framebuffer");
        use_set(c, v & 1);
        if (v < 2) CK(cudaStreamWaitEvent(c->own_stream, c->ev_fork, 0));
        FrameUniforms U;
        int rc = build_frame(c, views + 16 * v, projs + 16 * v, w, h, o, U);
        if (rc) return rc;
        rc = enqueue_frame(c, c->own_stream, fbs[v], U, flags | FRAME_NO_PDL);
        if (rc) return rc;
        if (v >= n_views - 2) {                      // join: the caller's stream continues after both sets drain
            CK(cudaEventRecord(c->ev_rendered[v & 1], c->own_stream));
            CK(cudaStreamWaitEvent(user, c->ev_rendered[v & 1], 0));
        }
    }
    return SPLAT_OK;        // the current set is the last view's: splat_frame_status / debug fetch describe that frame
}

int splat_frame_status(splat_ctx* c, splat_status* out) {
    if (!c || !out) ARG_FAIL("splat_frame_status: NULL argument");
    CK(cudaSetDevice(c->device));
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(c->ctl_host, c->ctl, sizeof(FrameCtl), cudaMemcpyDeviceToHost));
    std::memset(out, 0, sizeof(*out));
    out->n_splats = c->n;
    out->n_pairs = c->ctl_host->total_pairs;
    out->pair_capacity = c->capacity;
    out->overflow = c->ctl_host->overflow;
    out->sort_passes = 1;            // one shared-memory sort per tile list
    out->n_candidates = (unsigned long long)c->ctl_host->n_chunks * PARK_CHUNK; out->max_tile_pairs = c->ctl_host->max_tile_pairs; out->n_live_blocks = c->ctl_host->n_live_blocks;
    out->blend_warp_candidates = c->ctl_host->blend_warp_candidates;
    // every frame since the last look (a camera batch: all its views), not only the last one
    unsigned long long worst = 0;
    const int rc = check_overflow(c, true, &worst);
    if (worst) { out->overflow = 1; if (worst > out->n_pairs) out->n_pairs = worst; }
    out->pair_capacity = c->capacity;
    return rc == SPLAT_ERR_OVERFLOW ? SPLAT_OK : rc;      // the overflow is reported in *out; the buffers are grown
}

int splat_render_host(splat_ctx* c, void* fb_host, const float* view, const float* proj, int w, int h,
                      const splat_opts* o) {
    if (!c) ARG_FAIL("splat_render_host: ctx is NULL");
    if (!fb_host) ARG_FAIL("splat_render_host: framebuffer is NULL");
    CK(cudaSetDevice(c->device));
    use_set(c, 0);
    FrameUniforms U;
    int rc = build_frame(c, view, proj, w, h, o, U);
    if (rc) return rc;
    const size_t bytes = (size_t)w * h * 8;
    if (bytes > c->fb_scratch_bytes) {
        CK(cudaDeviceSynchronize());
        cudaFree(c->fb_scratch); c->fb_scratch = nullptr;
        CK(cudaMalloc(&c->fb_scratch, bytes));
        c->fb_scratch_bytes = bytes;
    }
    if (c->capacity == 0) { rc = ensure_pairs(c, 1ull << 20); if (rc) return rc; }
    const bool bgra = o && (o->flags & SPLAT_FLAG_HOST_BGRA8);
    if (bgra && (size_t)w * h * 4 > c->bgra_bytes) {
        CK(cudaDeviceSynchronize());
        for (int k = 0; k < FB_RING; ++k) { cudaFree(c->bgra_scratch[k]); c->bgra_scratch[k] = nullptr; CK(cudaMalloc(&c->bgra_scratch[k], (size_t)w * h * 4)); }
        c->bgra_bytes = (size_t)w * h * 4;
    }
    // rows outside the band are not rendered: only the band is copied back
    const size_t row0 = (size_t)U.row_begin * SPLAT_TILE, row1 = (size_t)U.row_end * SPLAT_TILE < (size_t)h ? (size_t)U.row_end * SPLAT_TILE : (size_t)h;
    for (int attempt = 0; attempt < 4; ++attempt) {
        rc = enqueue_frame(c, c->own_stream, c->fb_scratch, U, o ? o->flags : 0);
        if (rc) return rc;
        const uint32_t this_frame = c->frame_no;      // its status arrives in the mapped ring (written by the scatter kernel): no copy
        if (row1 > row0 && bgra) {
            launch_pack_bgra8((char*)c->fb_scratch + row0 * w * 8, (char*)c->bgra_scratch[0] + row0 * w * 4, (row1 - row0) * w, c->own_stream);
            CK(cudaMemcpyAsync((char*)fb_host + row0 * w * 4, (char*)c->bgra_scratch[0] + row0 * w * 4, (row1 - row0) * w * 4,
                               cudaMemcpyDeviceToHost, c->own_stream));
        } else if (row1 > row0)
            CK(cudaMemcpyAsync((char*)fb_host + row0 * w * 8, (char*)c->fb_scratch + row0 * w * 8, (row1 - row0) * w * 8,
                               cudaMemcpyDeviceToHost, c->own_stream));
        CK(cudaStreamSynchronize(c->own_stream));
        c->checked_frame = c->frame_no;          // this frame's status is handled right here
        const volatile FrameStatusSlot* sl = c->status_ring + (this_frame % STATUS_RING);
        if (sl->frame_no != this_frame) { g_err = "splat_render_host: the frame's status did not arrive"; return SPLAT_ERR_STATE; }
        if (!sl->overflow) return SPLAT_OK;
        const unsigned long long D = sl->total_slots;
        rc = ensure_pairs_all(c, D + D / 4 + 4096);
        if (rc) return rc;
    }
    g_err = "splat_render_host: pair buffers still overflow after regrowth";
    return SPLAT_ERR_OVERFLOW;
}

int splat_render_batch_host(splat_ctx* c, void* const* fb_host, const float* views, const float* projs, int n_views,
                            int w, int h, const splat_opts* o) {
    if (!c) ARG_FAIL("splat_render_batch_host: ctx is NULL");
    if (n_views < 0 || (n_views && (!fb_host || !views || !projs))) ARG_FAIL("splat_render_batch_host: bad arguments");
    if (n_views == 0) return SPLAT_OK;
    CK(cudaSetDevice(c->device));
    const size_t bytes = (size_t)w * h * 8;
    if (w <= 0 || h <= 0) ARG_FAIL("splat_render: bad viewport");
    if (bytes > c->fb_ring_bytes) {
        CK(cudaDeviceSynchronize());
        for (int k = 0; k < FB_RING; ++k) { cudaFree(c->fb_ring[k]); c->fb_ring[k] = nullptr; CK(cudaMalloc(&c->fb_ring[k], bytes)); }
        c->fb_ring_bytes = bytes;
    }
    if (n_views > STATUS_RING) {        // every view's status must still be in the ring when the batch has drained: take it in pieces
        for (int v0 = 0; v0 < n_views; v0 += STATUS_RING) {
            const int nv = n_views - v0 < STATUS_RING ? n_views - v0 : STATUS_RING;
            const int rc = splat_render_batch_host(c, fb_host + v0, views + 16 * v0, projs + 16 * v0, nv, w, h, o);
            if (rc) return rc;
        }
        return SPLAT_OK;
    }
    if (n_views > c->batch_flags_cap) {  // the views' frame numbers (plain host memory)
        std::free(c->batch_flags); c->batch_flags = nullptr;
        c->batch_flags = (unsigned int*)std::malloc((size_t)n_views * 4);
        if (!c->batch_flags) { c->batch_flags_cap = 0; g_err = "splat_render_batch_host: out of host memory"; return SPLAT_ERR_STATE; }
        c->batch_flags_cap = n_views;
    }
    use_set(c, 0);
    if (c->capacity == 0) { const int rc = ensure_pairs(c, 1ull << 20); if (rc) return rc; }
    const bool bgra = o && (o->flags & SPLAT_FLAG_HOST_BGRA8);
    if (bgra && (size_t)w * h * 4 > c->bgra_bytes) {
        CK(cudaDeviceSynchronize());
        for (int k = 0; k < FB_RING; ++k) { cudaFree(c->bgra_scratch[k]); c->bgra_scratch[k] = nullptr; CK(cudaMalloc(&c->bgra_scratch[k], (size_t)w * h * 4)); }
        c->bgra_bytes = (size_t)w * h * 4;
    }
    const size_t px_bytes = bgra ? 4 : 8;
    const bool two = n_views > 1 && two_frames_in_flight(c, w, h, o);
    if (two) { const int rc = ensure_alt(c); if (rc) return rc; }
    const int flags = o ? (o->flags & ~(SPLAT_FLAG_TIMING | SPLAT_FLAG_KEEP_UNSORTED)) : 0;
    void* const* buf = c->fb_ring;
    for (int attempt = 0; attempt < 4; ++attempt) {
        for (int v = 0; v < n_views; ++v) {
            if (!fb_host[v]) ARG_FAIL("splat_render_batch_host: NULL framebuffer");
            const int b = v % FB_RING;
            use_set(c, two ? (v & 1) : 0);                  // frame set and stream (two in flight); framebuffer ring slot b
            FrameUniforms U;
            int rc = build_frame(c, views + 16 * v, projs + 16 * v, w, h, o, U);
            if (rc) return rc;
            if (v >= FB_RING) CK(cudaStreamWaitEvent(c->own_stream, c->ev_copied[b], 0));     // slot b is free again
            rc = enqueue_frame(c, c->own_stream, buf[b], U, flags | (two ? FRAME_NO_PDL : 0));      // as splat_render_batch: two frames in flight fill each other's gaps
            if (rc) return rc;
            c->batch_flags[v] = c->frame_no;         // the view's status arrives in the mapped ring; a 24-byte copy here shared the
                                                     // copy engine with the previous frame's delivery and stalled this stream behind it
            const size_t row0 = (size_t)U.row_begin * SPLAT_TILE;
            const size_t row1 = (size_t)U.row_end * SPLAT_TILE < (size_t)h ? (size_t)U.row_end * SPLAT_TILE : (size_t)h;
            if (bgra && row1 > row0)
                launch_pack_bgra8((char*)buf[b] + row0 * w * 8, (char*)c->bgra_scratch[b] + row0 * w * 4, (row1 - row0) * w, c->own_stream);
            CK(cudaEventRecord(c->ev_rendered[b], c->own_stream));
            CK(cudaStreamWaitEvent(c->copy_stream, c->ev_rendered[b], 0));
            const char* src = bgra ? (const char*)c->bgra_scratch[b] : (const char*)buf[b];
            if (row1 > row0)
                CK(cudaMemcpyAsync((char*)fb_host[v] + row0 * w * px_bytes, src + row0 * w * px_bytes, (row1 - row0) * w * px_bytes,
                                   cudaMemcpyDeviceToHost, c->copy_stream));
            CK(cudaEventRecord(c->ev_copied[b], c->copy_stream));
        }
        CK(cudaStreamSynchronize(c->own_stream));
        if (c->alt_ready) CK(cudaStreamSynchronize(c->alt.own_stream));
        CK(cudaStreamSynchronize(c->copy_stream));
        c->checked_frame = c->frame_no;          // the views' status is handled right here
        unsigned long long worst = 0; bool over = false;
        for (int v = 0; v < n_views; ++v) {
            const volatile FrameStatusSlot* sl = c->status_ring + (c->batch_flags[v] % STATUS_RING);
            if (sl->frame_no != c->batch_flags[v]) { g_err = "splat_render_batch_host: a view's status did not arrive"; return SPLAT_ERR_STATE; }
            if (sl->overflow) { over = true; if (sl->total_slots > worst) worst = sl->total_slots; }
        }
        if (!over) return SPLAT_OK;
        const int rc = ensure_pairs_all(c, worst + worst / 4 + 4096);
        if (rc) return rc;
    }
    g_err = "splat_render_batch_host: pair buffers still overflow after regrowth";
    return SPLAT_ERR_OVERFLOW;
}

}  // extern "C"

// ------------------------------------------------------ debug / timing / sort
extern "C" {

// The pairs of the last frame as K1 parked them, put in the order the CPU restatement emits them: instance order, a
// splat's tiles row-major (= ascending tile id). The parking order itself depends on which warp got which chunk.
// tile[k], key[k] = depth field << 32 | instance index.
static cudaError_t fetch_parked(splat_ctx* c, std::vector<uint32_t>& tile, std::vector<unsigned long long>& key) {
    const size_t nch = c->ctl_host->n_chunks;
    tile.clear(); key.clear();
    if (!nch) return cudaSuccess;
    std::vector<unsigned long long> k(nch * PARK_CHUNK);
    std::vector<unsigned long long> t(nch * PARK_CHUNK); std::vector<uint32_t> fill(nch);
    cudaError_t e = cudaMemcpy(k.data(), c->park_key, k.size() * 8, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess) e = cudaMemcpy(t.data(), c->park_ts, t.size() * 8, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess) e = cudaMemcpy(fill.data(), c->chunk_fill, nch * 4, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) return e;
    std::vector<std::pair<unsigned long long, unsigned long long>> all;      // (instance << 32 | tile, key)
    for (size_t ch = 0; ch < nch; ++ch)
        for (uint32_t q = 0; q < fill[ch] && q < PARK_CHUNK; ++q) {
            const size_t g = ch * PARK_CHUNK + q;
            all.emplace_back(((k[g] & 0xFFFFFFFFull) << 32) | (t[g] >> 32), k[g]);
        }
    std::sort(all.begin(), all.end());
    tile.reserve(all.size()); key.reserve(all.size());
    for (const auto& a : all) { tile.push_back((uint32_t)a.first); key.push_back(a.second); }
    return cudaSuccess;
}

int splat_debug_fetch(splat_ctx* c, int what, void* dst, size_t* bytes) {
    if (!c || !bytes) ARG_FAIL("splat_debug_fetch: NULL argument");
    CK(cudaSetDevice(c->device));
    CK(cudaDeviceSynchronize());
    if (what == SPLAT_FETCH_DIGIT_HISTOGRAMS) {
        const size_t need = sizeof(c->sort_ctl->hist);
        if (!c->sort_ctl) { g_err = "splat_debug_fetch: no stand-alone sort has run"; return SPLAT_ERR_STATE; }
        if (!dst) { *bytes = need; return SPLAT_OK; }
        if (*bytes < need) { *bytes = need; ARG_FAIL("splat_debug_fetch: destination too small"); }
        CK(cudaMemcpy(dst, c->sort_ctl->hist, need, cudaMemcpyDeviceToHost));
        *bytes = need;
        return SPLAT_OK;
    }
    if (what == SPLAT_FETCH_INSTANCES || what == SPLAT_FETCH_SH_REST) {
        const bool sh = what == SPLAT_FETCH_SH_REST;
        if (sh && c->n && !c->sh) { g_err = "splat_debug_fetch: no SH coefficients uploaded"; return SPLAT_ERR_STATE; }
        const size_t need = (size_t)c->n * (sh ? 45 : 14) * 4;
        if (!dst) { *bytes = need; return SPLAT_OK; }
        if (*bytes < need) { *bytes = need; ARG_FAIL("splat_debug_fetch: destination too small"); }
        float* o = (float*)dst;
        std::vector<uint32_t> orig(c->n);            // plane entry i holds instance orig[i]
        if (c->n) CK(cudaMemcpy(orig.data(), c->orig, (size_t)c->n * 4, cudaMemcpyDeviceToHost));
        if (!sh && c->n) {
            std::vector<float> rot((size_t)c->n * 4), pos((size_t)c->n * 4), syz((size_t)c->n * 4), rgb((size_t)c->n * 4);
            CK(cudaMemcpy(rot.data(), c->rot, rot.size() * 4, cudaMemcpyDeviceToHost));
            CK(cudaMemcpy(pos.data(), c->pos_sx, pos.size() * 4, cudaMemcpyDeviceToHost));
            CK(cudaMemcpy(syz.data(), c->syz, syz.size() * 4, cudaMemcpyDeviceToHost));
            CK(cudaMemcpy(rgb.data(), c->base_rgb, rgb.size() * 4, cudaMemcpyDeviceToHost));
            for (size_t i = 0; i < c->n; ++i) {
                float* a = o + (size_t)orig[i] * 14;
                for (int k = 0; k < 4; ++k) a[k] = rot[i * 4 + k];
                a[4] = pos[i * 4 + 3]; a[5] = syz[i * 4 + 0]; a[6] = syz[i * 4 + 1];
                a[7] = pos[i * 4 + 0]; a[8] = pos[i * 4 + 1]; a[9] = pos[i * 4 + 2];
                a[10] = rgb[i * 4 + 0]; a[11] = rgb[i * 4 + 1]; a[12] = rgb[i * 4 + 2]; a[13] = syz[i * 4 + 2];
            }
        } else if (c->n) {
            // planes: float4[12][n_pad], coefficient-major; float index f of plane j = channel (4j+f)/15, coefficient (4j+f)%15
            std::vector<float> pl(c->n_pad * 48);
            CK(cudaMemcpy(pl.data(), c->sh, pl.size() * 4, cudaMemcpyDeviceToHost));
            for (size_t i = 0; i < c->n; ++i)
                for (int q = 0; q < 45; ++q) {
                    const int coef = q / 3, ch = q % 3;                   // plane order: coefficient-major, rgb interleaved
                    o[(size_t)orig[i] * 45 + ch * 15 + coef] = pl[((size_t)(q / 4) * c->n_pad + i) * 4 + (q % 4)];
                }
        }
        *bytes = need;
        return SPLAT_OK;
    }
    if (!c->have_frame) { g_err = "splat_debug_fetch: no frame rendered"; return SPLAT_ERR_STATE; }
    CK(cudaMemcpy(c->ctl_host, c->ctl, sizeof(FrameCtl), cudaMemcpyDeviceToHost));
    if (c->ctl_host->overflow) { g_err = "splat_debug_fetch: last frame overflowed the pair buffers"; return SPLAT_ERR_OVERFLOW; }
    const size_t D = (size_t)c->ctl_host->total_pairs;
    const size_t T = (size_t)c->lastU.tiles_x * c->lastU.tiles_y;
    const bool by_instance = c->lastU.mode != SPLAT_MODE_TILE;     // documented key = tile << 32 | instance index (D3)
    size_t need = 0;
    switch (what) {
        case SPLAT_FETCH_RECORDS:       need = (size_t)c->n * sizeof(splat_record); break;
        case SPLAT_FETCH_TILES_TOUCHED: need = (size_t)c->n * 4; break;
        case SPLAT_FETCH_OFFSETS:       need = (size_t)c->n * 4; break;
        case SPLAT_FETCH_KEYS_UNSORTED: need = D * 8; break;
        case SPLAT_FETCH_VALS_UNSORTED: need = D * 4; break;
        case SPLAT_FETCH_KEYS_SORTED:   need = D * 8; break;
        case SPLAT_FETCH_VALS_SORTED:   need = D * 4; break;
        case SPLAT_FETCH_TILE_RANGES:   need = T * 8; break;
        default: ARG_FAIL("splat_debug_fetch: unknown item");
    }
    if (!dst) { *bytes = need; return SPLAT_OK; }
    if (*bytes < need) { *bytes = need; ARG_FAIL("splat_debug_fetch: destination too small"); }
    std::vector<uint32_t> b(T + 1);
    CK(cudaMemcpy(b.data(), c->bounds, (T + 1) * 4, cudaMemcpyDeviceToHost));
    if (what == SPLAT_FETCH_TILE_RANGES) {
        uint32_t* r = (uint32_t*)dst;
        const size_t t_lo = (size_t)c->lastU.row_begin * c->lastU.tiles_x, t_hi = (size_t)c->lastU.row_end * c->lastU.tiles_x;
        for (size_t t = 0; t < T; ++t) {
            const bool any = t >= t_lo && t < t_hi && b[t] < b[t + 1];      // only the band's bounds were computed
            r[2 * t] = any ? b[t] : 0u; r[2 * t + 1] = any ? b[t + 1] : 0u;
        }
    } else if (what == SPLAT_FETCH_KEYS_SORTED || what == SPLAT_FETCH_VALS_SORTED) {
        // the list the pixels walk first: depth order (TILE), instance order (MLAB: the pairs carry no depth; LAYER50: the
        // arrival list), presented as tile << 32 | half-depth field, or | instance index for lists kept in instance order
        std::vector<unsigned long long> pr(D);
        std::vector<uint32_t> arr;
        if (D) CK(cudaMemcpy(pr.data(), c->pairs, D * 8, cudaMemcpyDeviceToHost));
        if (D && c->lastU.mode == SPLAT_MODE_LAYER50) { arr.resize(D); CK(cudaMemcpy(arr.data(), c->arrival, D * 4, cudaMemcpyDeviceToHost)); }
        size_t t = (size_t)c->lastU.row_begin * c->lastU.tiles_x;      // the band's first tile: its list starts at pair 0
        for (size_t i = 0; i < D; ++i) {
            while (t < T && b[t + 1] <= i) ++t;
            const uint32_t idx = arr.empty() ? (uint32_t)pr[i] : arr[i];
            if (what == SPLAT_FETCH_VALS_SORTED) ((uint32_t*)dst)[i] = idx;
            else ((uint64_t*)dst)[i] = ((uint64_t)t << 32) | (by_instance ? idx : (uint32_t)(pr[i] >> 32));
        }
    } else {
        std::vector<uint32_t> tile; std::vector<unsigned long long> key;
        CK(fetch_parked(c, tile, key));
        if (key.size() != D) { g_err = "splat_debug_fetch: parked pairs disagree with the tile counters"; return SPLAT_ERR_STATE; }
        if (what == SPLAT_FETCH_KEYS_UNSORTED)
            for (size_t i = 0; i < D; ++i) ((uint64_t*)dst)[i] = ((uint64_t)tile[i] << 32) | (by_instance ? (uint32_t)key[i] : (uint32_t)(key[i] >> 32));
        else if (what == SPLAT_FETCH_VALS_UNSORTED)
            for (size_t i = 0; i < D; ++i) ((uint32_t*)dst)[i] = (uint32_t)key[i];
        else {
            // pairs per splat (a splat whose tiles all failed the tile test counts as culled)
            std::vector<uint32_t> touched(c->n, 0u);
            for (size_t i = 0; i < D; ++i) touched[(uint32_t)key[i]]++;
            if (what == SPLAT_FETCH_TILES_TOUCHED) std::memcpy(dst, touched.data(), (size_t)c->n * 4);
            else if (what == SPLAT_FETCH_OFFSETS) {
                unsigned long long run = 0;
                for (uint32_t i = 0; i < c->n; ++i) { ((uint32_t*)dst)[i] = run > 0xFFFFFFFFull ? 0xFFFFFFFFu : (uint32_t)run; run += touched[i]; }
            } else if (c->n) {
                // device rows are 64 bytes (splat_record_row): present the 48-byte records, culled ones zeroed
                std::vector<splat_record_row> rows(c->n);
                CK(cudaMemcpy(rows.data(), c->rec, (size_t)c->n * sizeof(splat_record_row), cudaMemcpyDeviceToHost));
                splat_record* r = (splat_record*)dst;
                for (uint32_t i = 0; i < c->n; ++i) { if (touched[i]) r[i] = rows[i].r; else std::memset(r + i, 0, sizeof(splat_record)); }
            }
        }
    }
    *bytes = need;
    return SPLAT_OK;
}

int splat_pack_bgra8(splat_ctx* c, void* stream, const void* fb_half4, void* bgra8, int w, int h) {
    if (!c) ARG_FAIL("splat_pack_bgra8: ctx is NULL");
    if (w < 0 || h < 0 || ((w && h) && (!fb_half4 || !bgra8))) ARG_FAIL("splat_pack_bgra8: bad arguments");
    if (((uintptr_t)fb_half4 & 15u) || ((uintptr_t)bgra8 & 7u)) ARG_FAIL("splat_pack_bgra8: buffers must be 16- / 8-byte aligned");
    CK(cudaSetDevice(c->device));
    launch_pack_bgra8(fb_half4, bgra8, (size_t)w * h, (cudaStream_t)stream);
    CK(cudaGetLastError());
    return SPLAT_OK;
}

int splat_stage_times(splat_ctx* c, float ms[8]) {
    if (!c || !ms) ARG_FAIL("splat_stage_times: NULL argument");
    if (!c->have_frame || !c->timed) { g_err = "splat_stage_times: last frame had no SPLAT_FLAG_TIMING"; return SPLAT_ERR_STATE; }
    CK(cudaSetDevice(c->device));
    CK(cudaEventSynchronize(c->ev[5]));
    for (int k = 0; k < 8; ++k) ms[k] = 0.0f;
    for (int k = 0; k < 5; ++k) CK(cudaEventElapsedTime(&ms[k], c->ev[k], c->ev[k + 1]));
    CK(cudaEventElapsedTime(&ms[5], c->ev[0], c->ev[5]));
    return SPLAT_OK;
}

static int sort_pairs_impl(splat_ctx* c, void* stream, void* keys, uint32_t* vals, void* keys_alt, uint32_t* vals_alt,
                           uint64_t n, int begin_bit, int end_bit, int key_bytes, int* result_in_alt) {
    if (!c) ARG_FAIL("splat_sort_pairs: ctx is NULL");
    if (n && (!keys || !vals || !keys_alt || !vals_alt)) ARG_FAIL("splat_sort_pairs: NULL buffer");
    if (begin_bit < 0 || end_bit > key_bytes * 8 || begin_bit > end_bit) ARG_FAIL("splat_sort_pairs: bad bit range");
    if (n >= 0xFFFF0000ull) ARG_FAIL("splat_sort_pairs: n must be below 2^32");
    CK(cudaSetDevice(c->device));
    SortPlan plan{};
    for (int b = begin_bit; b < end_bit; b += 8) {
        const int bits = end_bit - b < 8 ? end_bit - b : 8;
        plan.shift[plan.n_pass] = b; plan.mask[plan.n_pass] = (1u << bits) - 1u; ++plan.n_pass;
    }
    for (int k = 0; k < SPLAT_MAX_SORT_PASSES; ++k) { plan.slot[k] = k; plan.ticket[k] = k; }
    if (result_in_alt) *result_in_alt = plan.n_pass & 1;
    if (n == 0 || plan.n_pass == 0) return SPLAT_OK;
    int rc = ensure_sort_status(c, n);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    c->sort_epoch = c->sort_epoch + 1;
    if (c->sort_epoch >= (1u << 25)) {      // epoch tags wrap: forget every lookback word of the previous cycle
        c->sort_epoch = 1;
        CK(cudaMemsetAsync(c->sort_status, 0, c->sort_status_tiles * 256 * 8, s));
    }
    CK(cudaMemsetAsync(c->sort_ctl, 0, sizeof(SortCtl), s));
    set_n_sort_kernel<<<1, 1, 0, s>>>(c->sort_ctl, (unsigned int)n);
    launch_sort_histogram(keys, key_bytes, n, c->sort_ctl, plan, c->sm_count, s);
    g_pdl_frame = true;
    launch_radix_sort(keys, vals, keys_alt, vals_alt, key_bytes, n, c->sort_ctl, c->sort_status,
                      c->sort_epoch * 16u + 1, plan, c->sm_count, false, s);
    CK(cudaGetLastError());
    return SPLAT_OK;
}

int splat_sort_pairs_u64_u32(splat_ctx* c, void* stream, uint64_t* keys, uint32_t* vals, uint64_t* keys_alt,
                             uint32_t* vals_alt, uint64_t n, int begin_bit, int end_bit, int* result_in_alt) {
    return sort_pairs_impl(c, stream, keys, vals, keys_alt, vals_alt, n, begin_bit, end_bit, 8, result_in_alt);
}

int splat_sort_pairs_u32_u32(splat_ctx* c, void* stream, uint32_t* keys, uint32_t* vals, uint32_t* keys_alt,
                             uint32_t* vals_alt, uint64_t n, int begin_bit, int end_bit, int* result_in_alt) {
    return sort_pairs_impl(c, stream, keys, vals, keys_alt, vals_alt, n, begin_bit, end_bit, 4, result_in_alt);
}

}  // extern "C"

// ------------------------------------------------------ multi-GPU single frame: tile-row bands + NCCL gather
// one context per device, each renders a band of tile rows (splat_opts.tile_row_begin / end) into a
// framebuffer of the full size; the path's one exchange step brings the bands together on rank `dst`. NCCL is taken from
// the calling process (the symbols are looked up at run time: a host that links NCCL, or PyTorch's bundled copy), so this
// library has no link-time dependency on it.
#include <dlfcn.h>
#include <nccl.h>

extern "C" {

int splat_band_rows(int tiles_y, int world, const double* row_weights, int32_t* bands) {
    if (tiles_y < 0 || world <= 0 || !bands) ARG_FAIL("splat_band_rows: bad arguments");
    if (!row_weights) {
        for (int r = 0; r < world; ++r) { bands[2 * r] = (int32_t)((long long)tiles_y * r / world); bands[2 * r + 1] = (int32_t)((long long)tiles_y * (r + 1) / world); }
        return SPLAT_OK;
    }
    // boundaries that minimise the heaviest band (binary search on the bottleneck), no rank left without a row while rows remain
    std::vector<double> wt(row_weights, row_weights + tiles_y);
    double sum = 0.0, mx = 0.0;
    for (double v : wt) { if (!(v >= 0.0)) ARG_FAIL("splat_band_rows: weights must be non-negative"); sum += v; }
    const double eps = 1e-9 * (sum > 1.0 ? sum : 1.0) / (tiles_y > 1 ? tiles_y : 1);
    sum = 0.0;
    for (double& v : wt) { v += eps; sum += v; if (v > mx) mx = v; }
    auto cuts = [&](double limit, std::vector<int32_t>& out) {
        out.clear();
        int start = 0; double acc = 0.0;
        for (int i = 0; i < tiles_y; ++i) {
            if (acc + wt[i] > limit && i > start) { out.push_back(start); out.push_back(i); start = i; acc = 0.0; }
            acc += wt[i];
        }
        out.push_back(start); out.push_back(tiles_y);
    };
    double lo = tiles_y ? mx : 0.0, hi = sum;
    std::vector<int32_t> b;
    for (int it = 0; it < 60; ++it) {
        const double mid = 0.5 * (lo + hi);
        cuts(mid, b);
        if ((int)b.size() / 2 <= world) hi = mid; else lo = mid;
    }
    cuts(hi, b);
    while ((int)b.size() / 2 < world) {          // more ranks than needed: split the widest band
        int k = 0;
        for (int j = 0; j < (int)b.size() / 2; ++j) if (b[2 * j + 1] - b[2 * j] > b[2 * k + 1] - b[2 * k]) k = j;
        const int32_t s0 = b[2 * k], e0 = b[2 * k + 1];
        if (e0 - s0 < 2) { b.push_back(tiles_y); b.push_back(tiles_y); continue; }
        const int32_t m = (s0 + e0) / 2;
        b[2 * k + 1] = m;
        b.insert(b.begin() + 2 * k + 2, {m, e0});
    }
    std::vector<std::pair<int32_t, int32_t>> srt;
    for (int j = 0; j < world; ++j) srt.emplace_back(b[2 * j], b[2 * j + 1]);
    std::sort(srt.begin(), srt.end());
    for (int j = 0; j < world; ++j) { bands[2 * j] = srt[j].first; bands[2 * j + 1] = srt[j].second; }
    return SPLAT_OK;
}

int splat_gather_bands_batch(void* nccl_comm, void* stream, void* const* fbs_dev, int n_fb, const int32_t* pixel_rows, int w,
                             int bytes_per_pixel, int rank, int world, int dst) {
    if (!nccl_comm || !pixel_rows || w <= 0 || bytes_per_pixel <= 0 || world <= 0 || rank < 0 || rank >= world || dst < 0 || dst >= world ||
        n_fb < 0 || (n_fb && !fbs_dev))
        ARG_FAIL("splat_gather_bands: bad arguments");
    for (int f = 0; f < n_fb; ++f) if (!fbs_dev[f]) ARG_FAIL("splat_gather_bands: NULL framebuffer");
    if (world == 1 || n_fb == 0) return SPLAT_OK;
    typedef ncclResult_t (*group_fn)(void);
    typedef ncclResult_t (*p2p_send_fn)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
    typedef ncclResult_t (*p2p_recv_fn)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
    static group_fn g_start = nullptr, g_end = nullptr; static p2p_send_fn g_send = nullptr; static p2p_recv_fn g_recv = nullptr;
    if (!g_start) {
        void* h = RTLD_DEFAULT;
        if (!dlsym(h, "ncclGroupStart")) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) { g_err = "splat_gather_bands: NCCL is not loaded in this process and libnccl.so.2 was not found"; return SPLAT_ERR_STATE; }
        g_end = (group_fn)dlsym(h, "ncclGroupEnd"); g_send = (p2p_send_fn)dlsym(h, "ncclSend"); g_recv = (p2p_recv_fn)dlsym(h, "ncclRecv");
        g_start = (group_fn)dlsym(h, "ncclGroupStart");
        if (!g_start || !g_end || !g_send || !g_recv) { g_start = nullptr; g_err = "splat_gather_bands: NCCL symbols not found"; return SPLAT_ERR_STATE; }
    }
    ncclComm_t comm = (ncclComm_t)nccl_comm; cudaStream_t s = (cudaStream_t)stream;
    const size_t row_bytes = (size_t)w * bytes_per_pixel;
    // ONE group for every framebuffer: NCCL fuses the sends / receives of a peer into one kernel launch
    ncclResult_t rc = g_start();
    for (int f = 0; f < n_fb && rc == ncclSuccess; ++f) {
        if (rank == dst) {
            for (int r = 0; r < world && rc == ncclSuccess; ++r) {
                const int32_t a = pixel_rows[2 * r], b = pixel_rows[2 * r + 1];
                if (r != dst && b > a) rc = g_recv((char*)fbs_dev[f] + (size_t)a * row_bytes, (size_t)(b - a) * row_bytes, ncclUint8, r, comm, s);
            }
        } else {
            const int32_t a = pixel_rows[2 * rank], b = pixel_rows[2 * rank + 1];
            if (b > a) rc = g_send((const char*)fbs_dev[f] + (size_t)a * row_bytes, (size_t)(b - a) * row_bytes, ncclUint8, dst, comm, s);
        }
    }
    const ncclResult_t rc2 = g_end();
    if (rc != ncclSuccess || rc2 != ncclSuccess) { g_err = "splat_gather_bands: NCCL send / recv failed"; return SPLAT_ERR_CUDA; }
    return SPLAT_OK;
}

int splat_gather_bands(void* nccl_comm, void* stream, void* fb_dev, const int32_t* pixel_rows, int w, int bytes_per_pixel,
                       int rank, int world, int dst) {
    if (!fb_dev) ARG_FAIL("splat_gather_bands: bad arguments");
    void* one[1] = {fb_dev};
    return splat_gather_bands_batch(nccl_comm, stream, one, 1, pixel_rows, w, bytes_per_pixel, rank, world, dst);
}

}  // extern "C"

// ------------------------------------------------------ multi-GPU single frame, fused exchange: peer framebuffers
// The gather as stores instead of a collective: the destination rank allocates its framebuffers here and exports them
// (CUDA IPC); every other rank opens them and passes the mapped pointer to splat_render as ITS framebuffer, so its blend
// kernel writes the band's pixels straight into the destination's memory over NVLink - no gather kernel, no SM time on the
// destination, the transfer spread over the blend's run time. What remains to exchange is a completion signal (any
// stream-ordered cross-GPU synchronisation after the render, e.g. a 4-byte NCCL all-reduce).
extern "C" {

int splat_peer_alloc(size_t bytes, void** ptr, void* handle64) {
    if (!ptr || !handle64 || bytes == 0) ARG_FAIL("splat_peer_alloc: bad arguments");
    *ptr = nullptr;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "the exported handle is 64 bytes");
    void* p = nullptr;
    CK(cudaMalloc(&p, bytes));
    cudaIpcMemHandle_t h;
    const cudaError_t e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) { cudaFree(p); g_err = std::string("splat_peer_alloc: cudaIpcGetMemHandle: ") + cudaGetErrorString(e); return SPLAT_ERR_CUDA; }
    std::memcpy(handle64, &h, 64);
    CK(cudaMemset(p, 0, bytes));
    *ptr = p;
    return SPLAT_OK;
}

int splat_peer_free(void* ptr) {
    if (ptr) CK(cudaFree(ptr));
    return SPLAT_OK;
}

int splat_peer_open(const void* handle64, void** ptr) {
    if (!ptr || !handle64) ARG_FAIL("splat_peer_open: bad arguments");
    *ptr = nullptr;
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle64, 64);
    void* p = nullptr;
    const cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) { g_err = std::string("splat_peer_open: cudaIpcOpenMemHandle: ") + cudaGetErrorString(e); return SPLAT_ERR_CUDA; }
    *ptr = p;
    return SPLAT_OK;
}

int splat_peer_close(void* ptr) {
    if (ptr) CK(cudaIpcCloseMemHandle(ptr));
    return SPLAT_OK;
}

}  // extern "C"
