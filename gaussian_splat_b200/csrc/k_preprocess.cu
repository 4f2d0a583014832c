// K0 upload (AoS SplatInstance -> SoA planes) and K1 preprocess.
// K1 replaces vertex_main (shaders/gaussian_splat_tile.metal:85-157), run once per splat
// instead of six times, plus the clip / tile-bbox part of the fixed-function rasteriser.
#include "splat_math.cuh"

// AoS rows are 14 floats (src/instanced_splat_renderer.h:6-11): rot[4] scale[3] pos[3] color[4].
__global__ void __launch_bounds__(256) upload_soa_kernel(const float* __restrict__ aos, unsigned int n,
                                                         float4* __restrict__ rot, float4* __restrict__ pos_sx,
                                                         float4* __restrict__ syz_op_rg, uint16_t* __restrict__ bq)
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
}

void launch_upload_soa(const float* aos_dev, unsigned int n, float4* rot, float4* pos_sx, float4* syz_op_rg,
                       uint16_t* b, cudaStream_t s)
{
    if (n == 0) return;
    upload_soa_kernel<<<(n + 255) / 256, 256, 0, s>>>(aos_dev, n, rot, pos_sx, syz_op_rg, b);
}

__global__ void __launch_bounds__(256) preprocess_kernel(ScenePlanes sp, const __grid_constant__ FrameUniforms U,
                                                         splat_record* __restrict__ rec,
                                                         unsigned int* __restrict__ touched)
{
    const unsigned int i = blockIdx.x * 256u + threadIdx.x;
    if (i >= U.n) return;
    const float4 q = __ldg(sp.rot + i);
    const float4 p = __ldg(sp.pos_sx + i);
    const float4 t = __ldg(sp.syz_op_rg + i);
    const unsigned int bh = __ldg(sp.b + i);
    splat_record r;
    const unsigned int cnt = preprocess_one(q.x, q.y, q.z, q.w, p.w, t.x, t.y, p.x, p.y, p.z,
                                            __float_as_uint(t.w), bh, t.z, U, r);
    touched[i] = cnt;
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
