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
    rot[i] = make_float4(r[0], r[1], r[2], r[3]);
    pos_sx[i] = make_float4(r[7], r[8], r[9], r[4]);
    const unsigned int rg = (unsigned int)__half_as_ushort(__float2half_rn(r[10])) |
                            ((unsigned int)__half_as_ushort(__float2half_rn(r[11])) << 16);
    syz_op_rg[i] = make_float4(r[5], r[6], r[13], __uint_as_float(rg));
    bq[i] = __half_as_ushort(__float2half_rn(r[12]));
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

__global__ void __launch_bounds__(256) preprocess_kernel(ScenePlanes sp, const __grid_constant__ FrameUniforms U,
                                                         splat_record* __restrict__ rec,
                                                         unsigned int* __restrict__ touched)
{
    const unsigned int i = blockIdx.x * 256u + threadIdx.x;
    if (i >= U.n) return;
    const float4 p = __ldg(sp.pos_sx + i);
    const float4 t = __ldg(sp.syz_op_rg + i);
    if (U.band_bound > 0.0f) {
        // Tile-row band (multi-GPU): drop splats that provably cannot reach the band before the ~200-op
        // covariance pipeline. The quad's pixel half-height is at most 3*sqrt(trace(Cov2D)) and
        // trace(J C J^T) <= ||J||_F^2 * ||V3x3||_F^2 * max(scale)^2 (+2e-4 for the two epsilons), so the test
        // is conservative (2% + 1 px slack for rounding) and never changes a result: the full path would
        // have clamped such a rectangle to an empty row range.
        float vp[4], cl[4];
        mat4_point(U.V, p.x, p.y, p.z, vp);
        const float zf = -vp[2];
        if (!(zf >= 1e-4f)) { touched[i] = 0; return; }
        mat4_point(U.VP, p.x, p.y, p.z, cl);
        const float cy = (0.5f - 0.5f * (cl[1] * (1.0f / cl[3]))) * U.H;
        const float iz = 1.0f / zf, fx = U.P[0] * (U.W * 0.5f) * iz, fy = U.P[5] * (U.H * 0.5f) * iz;
        const float jx = fx * vp[0] * iz, jy = fy * vp[1] * iz;
        const float smax = fmaxf(p.w, fmaxf(t.x, t.y));
        const float tr = (fx * fx + fy * fy + jx * jx + jy * jy) * U.band_bound * (smax * smax) + 2e-4f;
        const float hb = 3.0f * sqrtf(tr) + 1.0f;
        if (!(cy + hb >= (float)(U.row_begin * 16) && cy - hb <= (float)(U.row_end * 16))) { touched[i] = 0; return; }
    }
    const float4 q = __ldg(sp.rot + i);
    const unsigned int bh = __ldg(sp.b + i);
    splat_record r;
    const unsigned int cnt = preprocess_one(q.x, q.y, q.z, q.w, p.w, t.x, t.y, p.x, p.y, p.z,
                                            __float_as_uint(t.w), bh, t.z, U, r);
    touched[i] = cnt;
    if (cnt && U.sh_degree > 0) {
        // D11: colour = max(0, loader colour + sum_{k>=1} basis_k(dir) * f_rest[k][channel]), then half
        const float dx = p.x - U.cam[0], dy = p.y - U.cam[1], dz = p.z - U.cam[2];
        const float dl = sqrtf((dx * dx + dy * dy) + dz * dz);
        float bs[15];
        sh_basis(dx / dl, dy / dl, dz / dl, U.sh_degree, bs);
        const float4 base = __ldg(sp.base_rgb + i);
        float col[3] = {base.x, base.y, base.z};
        const int n_coef = (U.sh_degree + 1) * (U.sh_degree + 1) - 1;      // 3, 8, 15
        const int n_vec = (3 * n_coef + 3) / 4;                            // 3, 6, 12 float4 planes
        for (int j = 0; j < n_vec; ++j) {
            const float4 v = __ldg(sp.sh + (size_t)j * sp.n_pad + i);
            const float c4[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int f = 4 * j + q;
                if (f < 3 * n_coef) col[f % 3] = col[f % 3] + bs[f / 3] * c4[q];
            }
        }
        r.rg = (unsigned int)__half_as_ushort(__float2half_rn(fmaxf(col[0], 0.0f))) |
               ((unsigned int)__half_as_ushort(__float2half_rn(fmaxf(col[1], 0.0f))) << 16);
        r.bd = (r.bd & 0xFFFF0000u) | (unsigned int)__half_as_ushort(__float2half_rn(fmaxf(col[2], 0.0f)));
    }
    if (cnt) {
        float4* dst = reinterpret_cast<float4*>(rec + i);
        dst[0] = make_float4(r.cx, r.cy, r.ux, r.uy);
        dst[1] = make_float4(r.vx, r.vy, r.opacity, __uint_as_float(r.rg));
        dst[2] = make_float4(__uint_as_float(r.bd), r.z_front, __uint_as_float(r.rect_lo), __uint_as_float(r.rect_hi));
    }
}

void launch_preprocess(const ScenePlanes& sp, const FrameUniforms& U, splat_record* rec, unsigned int* touched,
                       cudaStream_t s)
{
    if (U.n == 0) return;
    preprocess_kernel<<<(U.n + 255) / 256, 256, 0, s>>>(sp, U, rec, touched);
}
