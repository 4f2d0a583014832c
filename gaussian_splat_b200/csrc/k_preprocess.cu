// K0 upload (AoS SplatInstance -> SoA planes) and K1 preprocess.
// K1 replaces vertex_main (shaders/gaussian_splat_tile.metal:85-157), run once per splat
// instead of six times, plus the clip / tile-bbox part of the fixed-function rasteriser.
#include "splat_math.cuh"

// AoS rows are 14 floats (src/instanced_splat_renderer.h:6-11): rot[4] scale[3] pos[3] color[4].
__global__ void __launch_bounds__(256) upload_soa_kernel(const float* __restrict__ aos, unsigned int n,
                                                         float4* __restrict__ rot, float4* __restrict__ pos_sx,
                                                         float4* __restrict__ syz_op_rg, uint16_t* __restrict__ bq,
                                                         float4* __restrict__ base_rgb)
{
    // Stage 256 rows (256*14 floats) through shared memory so the AoS read is coalesced.
    __shared__ float s[256 * 14];
    const unsigned int base = blockIdx.x * 256u;
    const unsigned int rows = min(256u, n - base);
    const float* src = aos + (size_t)base * 14;
    for (unsigned int k = threadIdx.x; k < rows * 14; k += 256) s[k] = src[k];
    __syncthreads();
    if (threadIdx.x >= rows) return;
    const float* r = s + threadIdx.x * 14;   // stride 14 words: conflict-free enough (gcd(14,32)=2)
    const unsigned int i = base + threadIdx.x;
    const float4 v_rot = make_float4(r[0], r[1], r[2], r[3]), v_pos = make_float4(r[7], r[8], r[9], r[4]);
    const unsigned int rg = (unsigned int)__half_as_ushort(__float2half_rn(r[10])) |
                            ((unsigned int)__half_as_ushort(__float2half_rn(r[11])) << 16);
    const float4 v_syz = make_float4(r[5], r[6], r[13], __uint_as_float(rg));
    const unsigned int v_b = __half_as_ushort(__float2half_rn(r[12]));
    rot[i] = v_rot; pos_sx[i] = v_pos; syz_op_rg[i] = v_syz; bq[i] = (uint16_t)v_b;
    base_rgb[i] = make_float4(r[10], r[11], r[12], 0.0f);
}

// f_rest rows (45 floats, channel-major as src/ply_loader.cpp:76-81 stores them) -> 12 float4 planes,
// coefficient-major, staged through shared memory so both sides are coalesced.
__global__ void __launch_bounds__(256) upload_sh_kernel(const float* __restrict__ sh, unsigned int n, size_t n_pad,
                                                        float4* __restrict__ planes)
{
    __shared__ float s[256 * 45];
    const unsigned int base = blockIdx.x * 256u;
    const unsigned int rows = min(256u, n - base);
    const float* src = sh + (size_t)base * 45;
    for (unsigned int k = threadIdx.x; k < rows * 45; k += 256) s[k] = src[k];
    __syncthreads();
    if (threadIdx.x >= rows) return;
    const float* r = s + threadIdx.x * 45;     // stride 45 words: conflict-free (odd)
#pragma unroll
    for (int j = 0; j < 12; ++j) {
        float v[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int f = 4 * j + q;               // f = 3*k + channel
            v[q] = f < 45 ? r[(f % 3) * 15 + f / 3] : 0.0f;
        }
        planes[(size_t)j * n_pad + base + threadIdx.x] = make_float4(v[0], v[1], v[2], v[3]);
    }
}

void launch_upload_sh(const float* sh_rest45_dev, unsigned int n, size_t n_pad, float4* planes, cudaStream_t s)
{
    if (n == 0) return;
    upload_sh_kernel<<<(n + 255) / 256, 256, 0, s>>>(sh_rest45_dev, n, n_pad, planes);
}

void launch_upload_soa(const float* aos_dev, unsigned int n, float4* rot, float4* pos_sx, float4* syz_op_rg,
                       uint16_t* b, float4* base_rgb, cudaStream_t s)
{
    if (n == 0) return;
    upload_soa_kernel<<<(n + 255) / 256, 256, 0, s>>>(aos_dev, n, rot, pos_sx, syz_op_rg, b, base_rgb);
}

#define K1_TILE (SCAN_TILE_ITEMS / 2)      // splats per pair-count cell (two cells per duplication tile)

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

// K1 = preprocess_kernel + tile_scan_kernel.
// preprocess_kernel: a CTA owns 1024 splats. (1) Screening, 4 splats per thread with all loads issued first (position /
// scale planes read coalesced and kept in shared memory for phase 3); culled splats get a zero bin word. (2) The
// survivors' local indices are compacted in shared memory. (3) The exactly rounded ~600-instruction pipeline runs on
// dense warps of survivors. The CTA also leaves the candidate count (tile-rectangle area) of each of its two 512-splat
// cells for tile_scan_kernel, so K2 needs neither tickets nor a lookback chain. History: 512-splat CTAs left two warps
// working while six waited at the barrier; separate screening / exact kernels lost more to the extra launch and the
// list round trip than they won.
#define K1_ITEMS 4
__global__ void __launch_bounds__(256) preprocess_kernel(ScenePlanes sp, const __grid_constant__ FrameUniforms U,
                                                         splat_record_row* __restrict__ rec, unsigned long long* __restrict__ binw,
                                                         unsigned int* __restrict__ tile_sums)
{
    constexpr int K1_CTA_SPLATS = 256 * K1_ITEMS;
    static_assert(K1_CTA_SPLATS % K1_TILE == 0, "a CTA covers whole candidate-count cells");
    __shared__ unsigned short s_list[K1_CTA_SPLATS];
    __shared__ unsigned int s_count, s_sum[K1_CTA_SPLATS / K1_TILE];
    __shared__ float4 s_p[K1_CTA_SPLATS], s_t[K1_CTA_SPLATS];
    const unsigned int tid = threadIdx.x, lane = tid & 31;
    const unsigned int base = blockIdx.x * (unsigned int)K1_CTA_SPLATS;
    pdl_prologue();
    if (tid == 0) s_count = 0;
    if (tid < K1_CTA_SPLATS / K1_TILE) s_sum[tid] = 0;
    float4 p[K1_ITEMS], t[K1_ITEMS];
#pragma unroll
    for (int k = 0; k < K1_ITEMS; ++k) {
        const unsigned int i = base + k * 256u + tid;
        if (i < U.n) { p[k] = __ldg(sp.pos_sx + i); t[k] = __ldg(sp.syz_op_rg + i); }
    }
    __syncthreads();
    // ---- phase 1: screen, compact the survivors' local indices (one shared atomic per warp for all its items)
    unsigned int wm[K1_ITEMS], wtotal = 0;
#pragma unroll
    for (int k = 0; k < K1_ITEMS; ++k) {
        const unsigned int li = k * 256u + tid, i = base + li;
        float4 ps = p[k];
        ps.w = fmaxf(p[k].w, fmaxf(t[k].x, t[k].y));      // x y z max(scale)
        const bool keep = i < U.n && may_be_visible(ps, U);
        if (i < U.n && !keep) binw[i] = 0ull;
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
    // ---- phase 2: the exactly rounded pipeline on dense warps of survivors
    const unsigned int n_keep = s_count;
    for (unsigned int j = tid; j < n_keep; j += 256u) {
        const unsigned int li = s_list[j], i = base + li;
        const float4 pp = s_p[li], tt = s_t[li];
        const float4 q = __ldg(sp.rot + i);
        const unsigned int bh = (unsigned int)__ldg(sp.b + i);
        splat_record r;
        const unsigned int cnt = preprocess_one(q.x, q.y, q.z, q.w, pp.w, tt.x, tt.y, pp.x, pp.y, pp.z,
                                                __float_as_uint(tt.w), bh, tt.z, U, r);
        binw[i] = cnt ? binw_pack(r.rect_lo, r.rect_hi, r.bd >> 16) : 0ull;
        if (cnt == 0) continue;
        atomicAdd(&s_sum[li / K1_TILE], cnt);       // fits: 512 splats x at most 65536 tiles each < 2^32
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
        float kx, ky;
        tile_ratios(r.ux, r.uy, r.vx, r.vy, kx, ky);         // what K2's per-tile test needs beside the record: the row's padding
        float4* dst = reinterpret_cast<float4*>(rec + i);
        dst[0] = make_float4(r.cx, r.cy, r.ux, r.uy);
        dst[1] = make_float4(r.vx, r.vy, r.opacity, __uint_as_float(r.rg));
        dst[2] = make_float4(__uint_as_float(r.bd), r.z_front, __uint_as_float(r.rect_lo), __uint_as_float(r.rect_hi));
        dst[3] = make_float4(kx, ky, 0.0f, 0.0f);           // both 32-byte sectors of the row written whole
    }
    __syncthreads();
    if (tid < K1_CTA_SPLATS / K1_TILE && base + tid * K1_TILE < U.n) tile_sums[blockIdx.x * (K1_CTA_SPLATS / K1_TILE) + tid] = s_sum[tid];
}

__global__ void __launch_bounds__(1024) tile_scan_kernel(const unsigned int* __restrict__ tile_sums, unsigned int* __restrict__ scan_prefix,
                                                         unsigned int n, unsigned long long capacity, FrameCtl* ctl)
{
    __shared__ unsigned long long s_wsum[32];
    pdl_prologue();
    const unsigned int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned int n_cells = (n + K1_TILE - 1) / K1_TILE;
    const unsigned int n_scan = (n + SCAN_TILE_ITEMS - 1) / SCAN_TILE_ITEMS;
    const unsigned int per = (n_scan + 1023u) / 1024u;          // contiguous chunk per thread: all its loads in flight at once
    const unsigned int j0 = min(tid * per, n_scan), j1 = min(j0 + per, n_scan);
    unsigned long long part = 0;
    for (unsigned int j = j0; j < j1; ++j) {
        part += tile_sums[2 * j];
        if (2 * j + 1 < n_cells) part += tile_sums[2 * j + 1];
    }
    unsigned long long incl = part;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const unsigned long long x = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= (unsigned)d) incl += x; }
    if (lane == 31) s_wsum[warp] = incl;
    __syncthreads();
    unsigned long long run = incl - part, total = 0;
    for (int k = 0; k < 32; ++k) { const unsigned long long x = s_wsum[k]; if (k < (int)warp) run += x; total += x; }
    for (unsigned int j = j0; j < j1; ++j) {
        scan_prefix[j] = run > 0xFFFFFFFFull ? 0xFFFFFFFFu : (unsigned int)run;     // saturating, as the offsets are
        run += tile_sums[2 * j];
        if (2 * j + 1 < n_cells) run += tile_sums[2 * j + 1];
    }
    if (tid == 0) {
        const bool over = total > capacity;
        ctl->total_cand = total;
        ctl->n_cand = over ? (unsigned int)capacity : (unsigned int)total;
        ctl->overflow = over ? 1u : 0u;
    }
}

void launch_preprocess(const ScenePlanes& sp, const FrameUniforms& U, splat_record_row* rec, unsigned long long* binw,
                       unsigned int* tile_sums, unsigned int* scan_prefix, unsigned long long capacity, FrameCtl* ctl, cudaStream_t s)
{
    if (U.n == 0) return;
    launch_pdl(preprocess_kernel, dim3((U.n + 1023u) / 1024u), dim3(256), 0, s, sp, U, rec, binw, tile_sums);
    launch_pdl(tile_scan_kernel, dim3(1), dim3(1024), 0, s, (const unsigned int*)tile_sums, scan_prefix, U.n, capacity, ctl);
}
