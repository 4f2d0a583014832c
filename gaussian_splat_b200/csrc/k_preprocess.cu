// K0 upload (AoS SplatInstance -> spatially ordered SoA planes + block bounds) and K1 preprocess + binning candidates.
// K1 replaces vertex_main (shaders/gaussian_splat_tile.metal:85-157), run once per splat
// instead of six times, plus the clip / tile-bbox part of the fixed-function rasteriser.
#include "splat_math.cuh"

// ------------------------------------------------------------------ K0: scene order
// The scene is stored in Morton order of the splat positions (63-bit codes, 21 bits per axis over the bounding box of
// the finite positions), SCENE_BLOCK (64) consecutive splats = one block with an axis-aligned bounding box and its largest scale.
// A frame first culls whole blocks against the view frustum / tile-row band (block_cull_kernel), so K1 touches only
// the splats of live blocks: a band of eight of the 6M / 4K frame keeps about a fifth of them instead of screening all
// 6M on every rank. Results do not depend on the order: every pair carries the ORIGINAL instance index (orig[]) and
// the tile lists are sorted on (depth, instance index), a total order.
__device__ __forceinline__ unsigned int float_ordered(float f) { const unsigned int u = __float_as_uint(f); return u ^ ((u >> 31) ? 0xFFFFFFFFu : 0x80000000u); }
__global__ void __launch_bounds__(256) scene_aabb_kernel(const float* __restrict__ aos, unsigned int n, unsigned int* __restrict__ mm)
{
    // mm[0..2] = min, mm[3..5] = max of the finite positions, as order-preserving unsigned keys
    unsigned int lo[3] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu}, hi[3] = {0u, 0u, 0u};
    for (unsigned int i = blockIdx.x * 256u + threadIdx.x; i < n; i += gridDim.x * 256u)
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const float v = aos[(size_t)i * 14 + 7 + a];
            if (isfinite(v)) { const unsigned int k = float_ordered(v); lo[a] = min(lo[a], k); hi[a] = max(hi[a], k); }
        }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) { lo[a] = min(lo[a], __shfl_xor_sync(0xFFFFFFFFu, lo[a], d)); hi[a] = max(hi[a], __shfl_xor_sync(0xFFFFFFFFu, hi[a], d)); }
        if ((threadIdx.x & 31) == 0) { atomicMin(mm + a, lo[a]); atomicMax(mm + 3 + a, hi[a]); }
    }
}
__device__ __forceinline__ unsigned long long spread21(unsigned int v) {      // bit k of v -> bit 3k
    unsigned long long x = v & 0x1FFFFFull;
    x = (x | (x << 32)) & 0x1F00000000FFFFull;
    x = (x | (x << 16)) & 0x1F0000FF0000FFull;
    x = (x | (x << 8)) & 0x100F00F00F00F00Full;
    x = (x | (x << 4)) & 0x10C30C30C30C30C3ull;
    x = (x | (x << 2)) & 0x1249249249249249ull;
    return x;
}
__global__ void __launch_bounds__(256) morton_kernel(const float* __restrict__ aos, unsigned int n, float3 lo, float3 scale,
                                                     unsigned long long* __restrict__ keys, unsigned int* __restrict__ vals)
{
    const unsigned int i = blockIdx.x * 256u + threadIdx.x;
    if (i >= n) return;
    const float x = aos[(size_t)i * 14 + 7], y = aos[(size_t)i * 14 + 8], z = aos[(size_t)i * 14 + 9];
    unsigned long long k = 0x7FFFFFFFFFFFFFFFull;          // non-finite positions (never visible) go last
    if (isfinite(x) && isfinite(y) && isfinite(z)) {
        const unsigned int qx = (unsigned int)fminf(fmaxf((x - lo.x) * scale.x, 0.0f), 2097151.0f);
        const unsigned int qy = (unsigned int)fminf(fmaxf((y - lo.y) * scale.y, 0.0f), 2097151.0f);
        const unsigned int qz = (unsigned int)fminf(fmaxf((z - lo.z) * scale.z, 0.0f), 2097151.0f);
        k = spread21(qx) | (spread21(qy) << 1) | (spread21(qz) << 2);
    }
    keys[i] = k; vals[i] = i;
}
void launch_scene_aabb(const float* aos_dev, unsigned int n, unsigned int* mm6_dev, cudaStream_t s)
{
    const unsigned int init[6] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0u, 0u, 0u};
    cudaMemcpyAsync(mm6_dev, init, sizeof(init), cudaMemcpyHostToDevice, s);
    if (n) scene_aabb_kernel<<<min((n + 255u) / 256u, 1184u), 256, 0, s>>>(aos_dev, n, mm6_dev);
}
void launch_morton(const float* aos_dev, unsigned int n, const float lo[3], const float hi[3], unsigned long long* keys, unsigned int* vals, cudaStream_t s)
{
    if (n == 0) return;
    float3 l = make_float3(lo[0], lo[1], lo[2]), sc;
    sc.x = hi[0] > lo[0] ? 2097151.0f / (hi[0] - lo[0]) : 0.0f;
    sc.y = hi[1] > lo[1] ? 2097151.0f / (hi[1] - lo[1]) : 0.0f;
    sc.z = hi[2] > lo[2] ? 2097151.0f / (hi[2] - lo[2]) : 0.0f;
    morton_kernel<<<(n + 255) / 256, 256, 0, s>>>(aos_dev, n, l, sc, keys, vals);
}

// AoS rows are 14 floats (src/instanced_splat_renderer.h:6-11): rot[4] scale[3] pos[3] color[4]. Plane entry i holds
// instance perm[i] (perm = the Morton order; NULL = identity).
__global__ void __launch_bounds__(256) upload_soa_kernel(const float* __restrict__ aos, const unsigned int* __restrict__ perm, unsigned int n,
                                                         float4* __restrict__ rot, float4* __restrict__ pos_sx,
                                                         float4* __restrict__ syz_op_rg, uint16_t* __restrict__ bq,
                                                         float4* __restrict__ base_rgb, unsigned int* __restrict__ orig)
{
    const unsigned int i = blockIdx.x * 256u + threadIdx.x;
    if (i >= n) return;
    const unsigned int src = perm ? perm[i] : i;
    const float2* r2 = reinterpret_cast<const float2*>(aos + (size_t)src * 14);      // rows are 56 bytes: 8-byte aligned
    float r[14];
#pragma unroll
    for (int k = 0; k < 7; ++k) { const float2 v = __ldg(r2 + k); r[2 * k] = v.x; r[2 * k + 1] = v.y; }
    const unsigned int rg = (unsigned int)__half_as_ushort(__float2half_rn(r[10])) |
                            ((unsigned int)__half_as_ushort(__float2half_rn(r[11])) << 16);
    rot[i] = make_float4(r[0], r[1], r[2], r[3]);
    pos_sx[i] = make_float4(r[7], r[8], r[9], r[4]);
    syz_op_rg[i] = make_float4(r[5], r[6], r[13], __uint_as_float(rg));
    bq[i] = __half_as_ushort(__float2half_rn(r[12]));
    base_rgb[i] = make_float4(r[10], r[11], r[12], 0.0f);
    orig[i] = src;
}

// Bounds of every block of SCENE_BLOCK consecutive (Morton-ordered) splats: lo = (min xyz of the finite positions, largest
// scale), hi = (max xyz, -). One warp per block. A block without a finite position keeps lo > hi (never visible).
__global__ void __launch_bounds__(256) block_bounds_kernel(const float4* __restrict__ pos_sx, const float4* __restrict__ syz_op_rg,
                                                           unsigned int n, unsigned int n_blocks, float4* __restrict__ blk_lo, float4* __restrict__ blk_hi)
{
    const unsigned int blk = blockIdx.x * 8u + (threadIdx.x >> 5), lane = threadIdx.x & 31u;
    if (blk >= n_blocks) return;
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY}, sm = 0.0f;
    bool wild = false;          // a non-finite or negative scale: the screening bound cannot be trusted, keep the block
    for (unsigned int k = lane; k < SCENE_BLOCK; k += 32u) {
        const unsigned int i = blk * SCENE_BLOCK + k;
        if (i >= n) break;
        const float4 p = pos_sx[i], t = syz_op_rg[i];
        if (isfinite(p.x) && isfinite(p.y) && isfinite(p.z)) {
            lo[0] = fminf(lo[0], p.x); lo[1] = fminf(lo[1], p.y); lo[2] = fminf(lo[2], p.z);
            hi[0] = fmaxf(hi[0], p.x); hi[1] = fmaxf(hi[1], p.y); hi[2] = fmaxf(hi[2], p.z);
            const float m = fmaxf(fabsf(p.w), fmaxf(fabsf(t.x), fabsf(t.y)));
            if (!isfinite(m)) wild = true; else sm = fmaxf(sm, m);
        }
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
#pragma unroll
        for (int a = 0; a < 3; ++a) { lo[a] = fminf(lo[a], __shfl_xor_sync(0xFFFFFFFFu, lo[a], d)); hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xFFFFFFFFu, hi[a], d)); }
        sm = fmaxf(sm, __shfl_xor_sync(0xFFFFFFFFu, sm, d));
    }
    wild = __any_sync(0xFFFFFFFFu, wild);
    if (lane == 0) { blk_lo[blk] = make_float4(lo[0], lo[1], lo[2], wild ? INFINITY : sm); blk_hi[blk] = make_float4(hi[0], hi[1], hi[2], 0.0f); }
}

// f_rest rows (45 floats, channel-major as src/ply_loader.cpp:76-81 stores them) -> 12 float4 planes,
// coefficient-major; plane entry i holds instance perm[i].
__global__ void __launch_bounds__(256) upload_sh_kernel(const float* __restrict__ sh, const unsigned int* __restrict__ perm, unsigned int n, size_t n_pad,
                                                        float4* __restrict__ planes)
{
    const unsigned int i = blockIdx.x * 256u + threadIdx.x;
    if (i >= n) return;
    const float* r = sh + (size_t)(perm ? perm[i] : i) * 45;
#pragma unroll
    for (int j = 0; j < 12; ++j) {
        float v[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int f = 4 * j + q;               // f = 3*k + channel
            v[q] = f < 45 ? __ldg(r + (f % 3) * 15 + f / 3) : 0.0f;
        }
        planes[(size_t)j * n_pad + i] = make_float4(v[0], v[1], v[2], v[3]);
    }
}

void launch_upload_sh(const float* sh_rest45_dev, const unsigned int* perm, unsigned int n, size_t n_pad, float4* planes, cudaStream_t s)
{
    if (n == 0) return;
    upload_sh_kernel<<<(n + 255) / 256, 256, 0, s>>>(sh_rest45_dev, perm, n, n_pad, planes);
}

void launch_upload_soa(const float* aos_dev, const unsigned int* perm, unsigned int n, float4* rot, float4* pos_sx, float4* syz_op_rg,
                       uint16_t* b, float4* base_rgb, unsigned int* orig, float4* blk_lo, float4* blk_hi, cudaStream_t s)
{
    if (n == 0) return;
    upload_soa_kernel<<<(n + 255) / 256, 256, 0, s>>>(aos_dev, perm, n, rot, pos_sx, syz_op_rg, b, base_rgb, orig);
    const unsigned int nb = (n + SCENE_BLOCK - 1u) / SCENE_BLOCK;
    block_bounds_kernel<<<(nb + 7) / 8, 256, 0, s>>>(pos_sx, syz_op_rg, n, nb, blk_lo, blk_hi);
}

// Screening: can this splat possibly produce a tile (in the viewport and, with tile-row bands, in the
// band)? Fast approximate arithmetic (FMA, rcp.approx, sqrt.approx: ~70 instructions against ~600 for the
// exactly rounded pipeline) with 5% + 2 px slack, so it only ever says "no" when the exact path would too:
// the quad's pixel half-extent is at most 3*sqrt(trace(Cov2D)) and trace(J W S W^T J^T) <= ||J||_F^2 *
// lambda_max(W^T W) * max(scale)^2 (+2e-4 for the two epsilons), W the view matrix's 3x3 block; the host
// passes a Gershgorin bound on lambda_max (1 for a rigid view, where the Frobenius norm it replaced gave 3). Splats too close to the eye plane for the
// bound to be trusted are kept. Results never depend on it; it exists so that the exact path runs on
// warps full of survivors instead of on every warp that holds one (splat order is unrelated to screen
// position: at 1M/1080p 46% of the splats are culled, with tile-row bands up to 97%).
__device__ __forceinline__ float rcp_ftz(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float sqrt_ftz(float x) { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ bool may_be_visible(const float4 p, const FrameUniforms& U) {      // p = x y z max(scale)
    // straight-line: lanes that fail the depth guard compute garbage (inf / nan) that the final select discards;
    // single-instruction reciprocal and square root (flush-to-zero: the guard keeps their arguments above 1e-3 / 2e-4)
    const float vx = fmaf(U.V[0], p.x, fmaf(U.V[4], p.y, fmaf(U.V[8], p.z, U.V[12])));
    const float vy = fmaf(U.V[1], p.x, fmaf(U.V[5], p.y, fmaf(U.V[9], p.z, U.V[13])));
    const float zf = -fmaf(U.V[2], p.x, fmaf(U.V[6], p.y, fmaf(U.V[10], p.z, U.V[14])));
    const float cw = fmaf(U.VP[3], p.x, fmaf(U.VP[7], p.y, fmaf(U.VP[11], p.z, U.VP[15])));
    const bool trusted = zf > 1e-3f && cw > 1e-3f;
    const bool behind = zf < -1e-3f && cw < 0.0f;            // clearly behind the eye: culled by zFront < 1e-4
    const float rw = rcp_ftz(cw), iz = rcp_ftz(zf);
    const float ndcz = fmaf(U.VP[2], p.x, fmaf(U.VP[6], p.y, fmaf(U.VP[10], p.z, U.VP[14]))) * rw;
    const float cx = fmaf(fmaf(U.VP[0], p.x, fmaf(U.VP[4], p.y, fmaf(U.VP[8], p.z, U.VP[12]))) * rw, U.scr_hw, U.scr_hw);
    const float cy = fmaf(fmaf(U.VP[1], p.x, fmaf(U.VP[5], p.y, fmaf(U.VP[9], p.z, U.VP[13]))) * rw, -U.scr_hh, U.scr_hh);
    // trace bound = iz^2 smax^2 B (Fx^2 (1 + (vx iz)^2) + Fy^2 (1 + (vy iz)^2)); Fx^2 B and Fy^2 B come from the host
    const float ax = vx * iz, ay = vy * iz, si = p.w * iz;
    const float g = fmaf(U.scr_fx2b, fmaf(ax, ax, 1.0f), U.scr_fy2b * fmaf(ay, ay, 1.0f));
    const float hb = fmaf(3.15f, sqrt_ftz(fmaf(g, si * si, 2e-4f)), 2.0f);
    const bool in_view = ndcz >= -2e-3f && ndcz <= 1.002f && cx + hb >= 0.0f && cx - hb <= U.W &&
                         cy + hb >= U.scr_ylo && cy - hb <= U.scr_yhi;
    return trusted ? in_view : !behind;
}

// Block culling: the same conservative bound as may_be_visible, taken over a block's bounding box. View depth and
// clip w are affine in the position, so their extremes over the box are at its corners; with w > 0 on the whole box the
// projected centre and NDC depth are ratios of affine functions, whose extremes over a convex set are at its vertices
// too. The radius bound takes the block's largest scale at its smallest depth. Blocks that reach the eye plane are kept.
__device__ __forceinline__ bool block_may_be_visible(const float4 lo, const float4 hi, const FrameUniforms& U) {
    if (!(lo.x <= hi.x)) return false;            // no finite position in the block
    if (!(lo.w <= 3.0e38f)) return true;          // wild scales: not bounded here
    float zf_min = INFINITY, zf_max = -INFINITY, cw_min = INFINITY, cw_max = -INFINITY, ax = 0.0f, ay = 0.0f;
    float sx_min = INFINITY, sx_max = -INFINITY, sy_min = INFINITY, sy_max = -INFINITY, nz_min = INFINITY, nz_max = -INFINITY;
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        const float x = (c & 1) ? hi.x : lo.x, y = (c & 2) ? hi.y : lo.y, z = (c & 4) ? hi.z : lo.z;
        const float vx = fmaf(U.V[0], x, fmaf(U.V[4], y, fmaf(U.V[8], z, U.V[12])));
        const float vy = fmaf(U.V[1], x, fmaf(U.V[5], y, fmaf(U.V[9], z, U.V[13])));
        const float zf = -fmaf(U.V[2], x, fmaf(U.V[6], y, fmaf(U.V[10], z, U.V[14])));
        const float cw = fmaf(U.VP[3], x, fmaf(U.VP[7], y, fmaf(U.VP[11], z, U.VP[15])));
        const float rw = rcp_ftz(cw);
        const float nz = fmaf(U.VP[2], x, fmaf(U.VP[6], y, fmaf(U.VP[10], z, U.VP[14]))) * rw;
        const float sx = fmaf(fmaf(U.VP[0], x, fmaf(U.VP[4], y, fmaf(U.VP[8], z, U.VP[12]))) * rw, U.scr_hw, U.scr_hw);
        const float sy = fmaf(fmaf(U.VP[1], x, fmaf(U.VP[5], y, fmaf(U.VP[9], z, U.VP[13]))) * rw, -U.scr_hh, U.scr_hh);
        zf_min = fminf(zf_min, zf); zf_max = fmaxf(zf_max, zf); cw_min = fminf(cw_min, cw); cw_max = fmaxf(cw_max, cw);
        ax = fmaxf(ax, fabsf(vx)); ay = fmaxf(ay, fabsf(vy));
        sx_min = fminf(sx_min, sx); sx_max = fmaxf(sx_max, sx); sy_min = fminf(sy_min, sy); sy_max = fmaxf(sy_max, sy);
        nz_min = fminf(nz_min, nz); nz_max = fmaxf(nz_max, nz);
    }
    // rounding of the affine forms grows with the coordinates: the guards scale with them
    const float mag = fmaxf(fmaxf(fabsf(lo.x), fabsf(hi.x)), fmaxf(fmaxf(fabsf(lo.y), fabsf(hi.y)), fmaxf(fabsf(lo.z), fabsf(hi.z))));
    const float tol = 1e-3f + 1e-5f * mag;
    if (!(zf_min > tol && cw_min > tol)) return !(zf_max < -tol && cw_max < 0.0f);
    const float iz = rcp_ftz(zf_min), si = lo.w * iz;
    ax *= iz; ay *= iz;
    const float g = fmaf(U.scr_fx2b, fmaf(ax, ax, 1.0f), U.scr_fy2b * fmaf(ay, ay, 1.0f));
    const float hb = fmaf(3.15f, sqrt_ftz(fmaf(g, si * si, 2e-4f)), 3.0f) + 1e-4f * (U.W + U.H);
    return nz_max >= -2e-3f && nz_min <= 1.002f && sx_max + hb >= 0.0f && sx_min - hb <= U.W &&
           sy_max + hb >= U.scr_ylo && sy_min - hb <= U.scr_yhi;
}

__global__ void __launch_bounds__(256) block_cull_kernel(ScenePlanes sp, const __grid_constant__ FrameUniforms U, FrameCtl* ctl,
                                                         unsigned int* __restrict__ live)
{
    pdl_prologue();
    const unsigned int blk = blockIdx.x * 256u + threadIdx.x, lane = threadIdx.x & 31u;
    const bool keep = blk < sp.n_blocks && block_may_be_visible(__ldg(sp.blk_lo + blk), __ldg(sp.blk_hi + blk), U);
    const unsigned int m = __ballot_sync(0xFFFFFFFFu, keep);
    if (!m) return;
    unsigned int base = 0;
    if (lane == 0) base = atomicAdd(&ctl->n_live_blocks, (unsigned int)__popc(m));
    base = __shfl_sync(0xFFFFFFFFu, base, 0);
    if (keep) live[base + __popc(m & ((1u << lane) - 1u))] = blk;
}

// K1: a CTA owns 1024 splats = 1024 / SCENE_BLOCK live blocks. (1) Screening, 4 splats per thread with all loads issued first (position
// / scale planes read coalesced and kept in shared memory for phase 2). (2) The survivors' local indices are compacted in
// shared memory, so the exactly rounded ~600-instruction pipeline runs on dense warps of survivors (splat order is
// unrelated to screen position inside a block; without this every warp holding one survivor pays the exact path).
// (3) A warp takes 32 survivors at a time: record rows are written, then the warp expands its splats' tile rectangles
// cooperatively - 32 consecutive candidate (splat, tile) slots per step, each lane finding its owner among the warp's
// 32 rectangles with five shuffles on the running sums and fetching the owner's test parameters with twelve more, all
// from registers. A candidate that passes the D9' tile test draws its slot in its tile's list from the tile's counter and is
// parked (key; tile, slot) in the warp's current 256-slot chunk of the parking arrays; chunks are handed out by one global
// atomic each and their fill is recorded, so the scatter kernel reads only what was written.
// History: 512-splat CTAs left two warps working while six waited at the barrier; separate screening / exact kernels
// lost more to the extra launch and the list round trip than they won; round 1's bin words + one-CTA prefix scan +
// separate duplication kernel (binary search per 32 slots, record gathers, returning atomics) took 40 + 6.5 + 41 us at
// 1M / 1080p where this kernel takes the 40 us of the first alone.
#ifndef K1_ITEMS
#define K1_ITEMS 4
#endif
#ifndef K1_MIN_CTAS
#define K1_MIN_CTAS 1
#endif
// TAIL = false: the CTA takes unit blockIdx.x (one unit = 1024 splats = K1_BLOCKS live blocks) and leaves: the main launch, whose
// grid is sized from the live-block count of an EARLIER frame of the same shape (the host's hint; all units when it has none) -
// a band's frame otherwise launched five CTAs for every one that found a live block. TAIL = true: a small second launch whose
// CTAs stride over the units beyond the main grid; it finds none unless the hint was stale, so a stale hint costs time, never a
// result. Two instantiations because the looping form needs 113 registers (two resident CTAs) where the single-unit form needs 64.
template <bool TAIL>
__global__ void __launch_bounds__(256, K1_MIN_CTAS) preprocess_kernel(ScenePlanes sp, const __grid_constant__ FrameUniforms U,
                                                         splat_record_row* __restrict__ rec, FrameCtl* ctl,
                                                         const unsigned int* live, unsigned long long* __restrict__ park_key,
                                                         unsigned long long* __restrict__ park_ts, unsigned int* __restrict__ chunk_fill,
                                                         unsigned long long capacity, unsigned int first_unit)
{
    constexpr int K1_CTA_SPLATS = 256 * K1_ITEMS;
    __shared__ unsigned short s_list[K1_CTA_SPLATS];
    constexpr int K1_BLOCKS = K1_CTA_SPLATS / (int)SCENE_BLOCK;      // live blocks a CTA takes
    __shared__ unsigned int s_count, s_blk[K1_BLOCKS];
    __shared__ float4 s_p[K1_CTA_SPLATS], s_t[K1_CTA_SPLATS];
    pdl_prologue();
    const unsigned int n_live = __ldcg(&ctl->n_live_blocks);
    const unsigned int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (unsigned int unit = first_unit + blockIdx.x; unit * K1_BLOCKS < n_live; unit += gridDim.x) {
    if (tid == 0) s_count = 0;
    if (tid < K1_BLOCKS) s_blk[tid] = unit * K1_BLOCKS + tid < n_live ? __ldcg(live + unit * K1_BLOCKS + tid) : 0xFFFFFFFFu;
    __syncthreads();
    float4 p[K1_ITEMS], t[K1_ITEMS];
    unsigned int idx[K1_ITEMS];
#pragma unroll
    for (int k = 0; k < K1_ITEMS; ++k) {
        const unsigned int li = k * 256u + tid, b = s_blk[li / SCENE_BLOCK];
        idx[k] = b == 0xFFFFFFFFu ? 0xFFFFFFFFu : b * SCENE_BLOCK + (li % SCENE_BLOCK);
        if (idx[k] < U.n) { p[k] = __ldg(sp.pos_sx + idx[k]); t[k] = __ldg(sp.syz_op_rg + idx[k]); }
    }
    // ---- phase 1: screen, compact the survivors' local indices (one shared atomic per warp for all its items)
    unsigned int wm[K1_ITEMS], wtotal = 0;
#pragma unroll
    for (int k = 0; k < K1_ITEMS; ++k) {
        float4 ps = p[k];
        ps.w = fmaxf(p[k].w, fmaxf(t[k].x, t[k].y));      // x y z max(scale)
        const bool keep = idx[k] < U.n && may_be_visible(ps, U);
        wm[k] = __ballot_sync(0xFFFFFFFFu, keep);
        wtotal += __popc(wm[k]);
    }
    unsigned int wbase = 0;
    if (lane == 0 && wtotal) wbase = atomicAdd(&s_count, wtotal);
    wbase = __shfl_sync(0xFFFFFFFFu, wbase, 0);
#pragma unroll
    for (int k = 0; k < K1_ITEMS; ++k) {
        const unsigned int li = k * 256u + tid;
        if ((wm[k] >> lane) & 1u) {
            s_list[wbase + __popc(wm[k] & ((1u << lane) - 1u))] = (unsigned short)li;
            s_p[li] = p[k]; s_t[li] = t[k];
        }
        wbase += __popc(wm[k]);
    }
    __syncthreads();
    // ---- phase 2: the exactly rounded pipeline on dense warps of survivors, then the warp's candidates
    const unsigned int n_keep = s_count;
    unsigned int* const tile_count = ctl_tile_count(ctl);
    const bool by_instance = U.mode == SPLAT_MODE_MLAB;       // the only list MLAB walks is in instance order: no depth in the key
    unsigned int chunk = 0xFFFFFFFFu, used = PARK_CHUNK;      // the warp's current parking chunk (none yet) and its fill
    unsigned int pend_g = 0xFFFFFFFFu, pend_tile = 0, pend_slot = 0;      // (tile, slot) of the previous step, stored one step late
    for (unsigned int j0 = warp * 32u; j0 < n_keep; j0 += 256u) {
        const unsigned int j = j0 + lane;
        unsigned int cnt = 0, o = 0;
        splat_record r;
        float kx = 0.0f, ky = 0.0f;
        if (j < n_keep) {
            const unsigned int li = s_list[j], i = s_blk[li / SCENE_BLOCK] * SCENE_BLOCK + (li % SCENE_BLOCK);
            const float4 pp = s_p[li], tt = s_t[li];
            const float4 q = __ldg(sp.rot + i);
            const unsigned int bh = (unsigned int)__ldg(sp.b + i);
            o = __ldg(sp.orig + i);
            cnt = preprocess_one(q.x, q.y, q.z, q.w, pp.w, tt.x, tt.y, pp.x, pp.y, pp.z, __float_as_uint(tt.w), bh, tt.z, U, r);
            if (cnt) {
                if (U.sh_degree > 0) {
                    // D11: colour = max(0, loader colour + sum_{k>=1} basis_k(dir) * f_rest[k][channel]), then half
                    const float dx = pp.x - U.cam[0], dy = pp.y - U.cam[1], dz = pp.z - U.cam[2];
                    const float dl = sqrtf((dx * dx + dy * dy) + dz * dz);
                    float bs[15];
                    sh_basis(dx / dl, dy / dl, dz / dl, U.sh_degree, bs);
                    const float4 base_c = __ldg(sp.base_rgb + i);
                    float col[3] = {base_c.x, base_c.y, base_c.z};
                    const int n_coef = (U.sh_degree + 1) * (U.sh_degree + 1) - 1;      // 3, 8, 15
                    const int n_vec = (3 * n_coef + 3) / 4;                            // 3, 6, 12 float4 planes
                    for (int jv = 0; jv < n_vec; ++jv) {
                        const float4 v = __ldg(sp.sh + (size_t)jv * sp.n_pad + i);
                        const float c4[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                        for (int qq = 0; qq < 4; ++qq) {
                            const int f = 4 * jv + qq;
                            if (f < 3 * n_coef) col[f % 3] = col[f % 3] + bs[f / 3] * c4[qq];
                        }
                    }
                    r.rg = (unsigned int)__half_as_ushort(__float2half_rn(fmaxf(col[0], 0.0f))) |
                           ((unsigned int)__half_as_ushort(__float2half_rn(fmaxf(col[1], 0.0f))) << 16);
                    r.bd = (r.bd & 0xFFFF0000u) | (unsigned int)__half_as_ushort(__float2half_rn(fmaxf(col[2], 0.0f)));
                }
                tile_ratios(r.ux, r.uy, r.vx, r.vy, kx, ky);
                float4* dst = reinterpret_cast<float4*>(rec + o);
                dst[0] = make_float4(r.cx, r.cy, r.ux, r.uy);
                dst[1] = make_float4(r.vx, r.vy, r.opacity, __uint_as_float(r.rg));
                dst[2] = make_float4(__uint_as_float(r.bd), r.z_front, __uint_as_float(r.rect_lo), __uint_as_float(r.rect_hi));
                dst[3] = make_float4(kx, ky, 0.0f, 0.0f);           // both 32-byte sectors of the row written whole
            }
        }
        // ---- the warp's candidates: running sums of the 32 rectangle areas, 32 consecutive slots per step
        unsigned int incl = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const unsigned int a = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= (unsigned)d) incl += a; }
        const unsigned int total = __shfl_sync(0xFFFFFFFFu, incl, 31);
        if (total == 0) continue;
        const unsigned int excl = incl - cnt;
        const unsigned int wdt = cnt ? (r.rect_hi & 0xFFFFu) - (r.rect_lo & 0xFFFFu) + 1u : 1u;
        const float inv_w = 1.0f / (float)wdt;
        const unsigned int h = cnt ? r.bd >> 16 : 0u;
        const unsigned int df = by_instance ? 0u : (U.order == SPLAT_ORDER_REFERENCE ? (0xFFFFu - h) : h);   // half-depth field (D3)
        const unsigned int pk = wdt | (df << 16);
        for (unsigned int s0 = 0; s0 < total; s0 += 32u) {
            const unsigned int s = s0 + lane;
            unsigned int c = 0;         // owner = first lane whose running sum exceeds s
#pragma unroll
            for (int st = 16; st > 0; st >>= 1) {
                const unsigned int e = __shfl_sync(0xFFFFFFFFu, incl, c + st - 1);
                if (e <= s) c += st;
            }
            const bool valid = s < total;
            const unsigned int own = valid ? c : 31u;
            const float o_cx = __shfl_sync(0xFFFFFFFFu, r.cx, own), o_cy = __shfl_sync(0xFFFFFFFFu, r.cy, own);
            const float o_ux = __shfl_sync(0xFFFFFFFFu, r.ux, own), o_uy = __shfl_sync(0xFFFFFFFFu, r.uy, own);
            const float o_vx = __shfl_sync(0xFFFFFFFFu, r.vx, own), o_vy = __shfl_sync(0xFFFFFFFFu, r.vy, own);
            const float o_kx = __shfl_sync(0xFFFFFFFFu, kx, own), o_ky = __shfl_sync(0xFFFFFFFFu, ky, own);
            const unsigned int o_lo = __shfl_sync(0xFFFFFFFFu, r.rect_lo, own), o_pk = __shfl_sync(0xFFFFFFFFu, pk, own);
            const unsigned int o_ex = __shfl_sync(0xFFFFFFFFu, excl, own), o_id = __shfl_sync(0xFFFFFFFFu, o, own);
            const float o_iw = __shfl_sync(0xFFFFFFFFu, inv_w, own);
            bool pass = false;
            unsigned int tile = 0;
            if (valid) {
                const unsigned int k = s - o_ex, w_ = o_pk & 0xFFFFu;
                int ry = (int)((float)k * o_iw), rx = (int)k - ry * (int)w_;      // k < 2^24: exact in float; quotient off by one at most
                if (rx < 0) { --ry; rx += (int)w_; } else if (rx >= (int)w_) { ++ry; rx -= (int)w_; }
                const int tx = (int)(o_lo & 0xFFFFu) + rx, ty = (int)(o_lo >> 16) + ry;
                pass = tile_touch(o_cx, o_cy, o_ux, o_uy, o_vx, o_vy, o_kx, o_ky, tx, ty, U.w, U.h);
                tile = (unsigned int)(ty * U.tiles_x + tx);
            }
            // the pair's slot in its tile's list: a returning atomic whose result is only stored - one step later, so its
            // round trip to L2 hides behind the next step's shuffles and tile test
            const unsigned int slot = pass ? atomicAdd(tile_count + tile, 1u) : 0u;
            const unsigned int m = __ballot_sync(0xFFFFFFFFu, pass);
            if (pend_g != 0xFFFFFFFFu) { park_ts[pend_g] = ((unsigned long long)pend_tile << 32) | pend_slot; pend_g = 0xFFFFFFFFu; }
            if (!m) continue;
            const unsigned int np = __popc(m);
            if (used + np > PARK_CHUNK) {           // next chunk: the fill of the old one is final
                if (chunk != 0xFFFFFFFFu && (unsigned long long)(chunk + 1u) * PARK_CHUNK <= capacity && lane == 0) chunk_fill[chunk] = used;
                if (lane == 0) chunk = atomicAdd(&ctl->n_chunks, 1u);
                chunk = __shfl_sync(0xFFFFFFFFu, chunk, 0);
                used = 0;
            }
            if ((unsigned long long)(chunk + 1u) * PARK_CHUNK <= capacity) {
                if (pass) {
                    const unsigned int g = chunk * PARK_CHUNK + used + __popc(m & ((1u << lane) - 1u));
                    park_key[g] = ((unsigned long long)(o_pk >> 16) << 32) | o_id;
                    pend_g = g; pend_tile = tile; pend_slot = slot;
                }
            } else if (lane == 0) ctl->overflow = 1u;       // the parking arrays are full: the frame builds no lists
            used += np;
        }
    }
    if (pend_g != 0xFFFFFFFFu) park_ts[pend_g] = ((unsigned long long)pend_tile << 32) | pend_slot;
    if (chunk != 0xFFFFFFFFu && (unsigned long long)(chunk + 1u) * PARK_CHUNK <= capacity && lane == 0) chunk_fill[chunk] = used;
    if (!TAIL) return;          // the main launch: one unit per CTA
    __syncthreads();            // the shared arrays are reused by the next unit
    }
}

void launch_preprocess(const ScenePlanes& sp, const FrameUniforms& U, splat_record_row* rec, FrameCtl* ctl, unsigned int* live,
                       unsigned long long* park_key, unsigned long long* park_ts, unsigned int* chunk_fill, unsigned long long capacity,
                       unsigned int live_blocks_hint, cudaStream_t s)
{
    if (U.n == 0) return;
    launch_pdl(block_cull_kernel, dim3((sp.n_blocks + 255u) / 256u), dim3(256), 0, s, sp, U, ctl, live);
    const unsigned int per_cta = 256u * K1_ITEMS / SCENE_BLOCK, all_units = (sp.n_blocks + per_cta - 1) / per_cta;
    unsigned int grid = all_units;
    if (live_blocks_hint) {     // 1/8 + 16 units of slack over the hint; never fewer CTAs than fill the machine once
        const unsigned int want = (live_blocks_hint + per_cta - 1) / per_cta;
        grid = want + want / 8u + 16u;
        if (grid < 592u) grid = 592u;
        if (grid > all_units) grid = all_units;
    }
    launch_pdl(preprocess_kernel<false>, dim3(grid), dim3(256), 0, s, sp, U, rec, ctl,
               (const unsigned int*)live, park_key, park_ts, chunk_fill, capacity, 0u);
    if (grid < all_units)
        launch_pdl(preprocess_kernel<true>, dim3(148), dim3(256), 0, s, sp, U, rec, ctl,
                   (const unsigned int*)live, park_key, park_ts, chunk_fill, capacity, grid);
}
