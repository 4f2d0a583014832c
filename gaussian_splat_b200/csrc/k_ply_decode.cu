// GPU-side PLY vertex decode: raw rows of 4-byte little-endian floats ->
// activated, cropped, order-preserving compacted SplatInstance AoS (+ the 45 f_rest floats).
// Restates, per row, src/ply_loader.cpp:91-146 (property -> PointData slot, sigmoid on opacity, exp on
// scale_*, DC -> colour with the all-zero quirk :133-139) and the constructor's crop and field copy
// src/instanced_splat_renderer.mm:360-388. The CPU loader (gaussian_splat_b200/host/ply_loader.cpp,
// itself pinned bit-for-bit to the reference's) is what the parity tests check this kernel against. NaN inputs
// give NaN outputs, but the device returns the canonical NaN where x86 keeps the operand's sign and payload.
#include "splat_internal.cuh"

#define PLY_ROWS 128           // rows per CTA, one thread per row
#define PLY_POINT_FLOATS 62    // PointData, src/ply_loader.h:7-17

// expf as the host C library computes it (glibc >= 2.27 / ARM optimized-routines expf: double arithmetic,
// 2^(k/32) table, degree-3 polynomial; |error| < 0.502 ulp), so std::exp(float) in the CPU loader and
// this kernel agree bit for bit. Every operation is a separately rounded double op, as in the C source.
__device__ __constant__ unsigned long long c_exp2f_tab[32] = {
    0x3ff0000000000000ull, 0x3fefd9b0d3158574ull, 0x3fefb5586cf9890full, 0x3fef9301d0125b51ull,
    0x3fef72b83c7d517bull, 0x3fef54873168b9aaull, 0x3fef387a6e756238ull, 0x3fef1e9df51fdee1ull,
    0x3fef06fe0a31b715ull, 0x3feef1a7373aa9cbull, 0x3feedea64c123422ull, 0x3feece086061892dull,
    0x3feebfdad5362a27ull, 0x3feeb42b569d4f82ull, 0x3feeab07dd485429ull, 0x3feea47eb03a5585ull,
    0x3feea09e667f3bcdull, 0x3fee9f75e8ec5f74ull, 0x3feea11473eb0187ull, 0x3feea589994cce13ull,
    0x3feeace5422aa0dbull, 0x3feeb737b0cdc5e5ull, 0x3feec49182a3f090ull, 0x3feed503b23e255dull,
    0x3feee89f995ad3adull, 0x3feeff76f2fb5e47ull, 0x3fef199bdd85529cull, 0x3fef3720dcef9069ull,
    0x3fef5818dcfba487ull, 0x3fef7c97337b9b5full, 0x3fefa4afa2a490daull, 0x3fefd0765b6e4540ull};

__device__ __forceinline__ float expf_libm(float x) {
    if (x != x) return x + x;
    if (x > 88.72283172607421875f) return __int_as_float(0x7F800000);       // > 0x1.62e42ep6f: overflow
    if (x < -103.97207641601562500f) return 0.0f;                           // < -0x1.9fe368p6f: underflow (also -inf)
    const double N = 32.0;
    const double inv_ln2_n = 1.4426950408889634 * N;                        // 0x1.71547652b82fep+0 * N
    const double shift = 6755399441055744.0;                                // 0x1.8p+52
    const double c0 = 0.05550361559341535 / N / N / N;                      // 0x1.c6af84b912394p-5
    const double c1 = 0.2402284522445722 / N / N;                           // 0x1.ebfce50fac4f3p-3
    const double c2 = 0.6931471806916203 / N;                               // 0x1.62e42ff0c52d6p-1
    const double z = __dmul_rn(inv_ln2_n, (double)x);
    double kd = __dadd_rn(z, shift);
    const unsigned long long ki = (unsigned long long)__double_as_longlong(kd);
    kd = __dsub_rn(kd, shift);
    const double r = __dsub_rn(z, kd);
    const unsigned long long t = c_exp2f_tab[ki & 31ull] + (ki << 47);
    const double s = __longlong_as_double((long long)t);
    const double zz = __dadd_rn(__dmul_rn(c0, r), c1);
    const double r2 = __dmul_rn(r, r);
    double y = __dadd_rn(__dmul_rn(c2, r), 1.0);
    y = __dadd_rn(__dmul_rn(zz, r2), y);
    y = __dmul_rn(y, s);
    return __double2float_rn(y);
}

__device__ __forceinline__ bool inside_crop(float x, float y, float z, float crop) {
    // !(fabs(x) < r && ...) drops NaNs too, as the reference's comparison does (instanced_splat_renderer.mm:363-365)
    return crop <= 0.0f || (fabsf(x) < crop && fabsf(y) < crop && fabsf(z) < crop);
}

// Pass 1: kept rows per CTA of PLY_ROWS rows (only x, y, z are touched).
__global__ void __launch_bounds__(PLY_ROWS) ply_count_kernel(const float* __restrict__ rows, unsigned int n_rows, unsigned int stride,
                                                             int jx, int jy, int jz, float crop, unsigned int* __restrict__ block_counts)
{
    const unsigned int i = blockIdx.x * PLY_ROWS + threadIdx.x;
    bool keep = false;
    if (i < n_rows) {
        const float* r = rows + (size_t)i * stride;
        keep = inside_crop(jx >= 0 ? __ldg(r + jx) : 0.0f, jy >= 0 ? __ldg(r + jy) : 0.0f, jz >= 0 ? __ldg(r + jz) : 0.0f, crop);
    }
    const int c = __syncthreads_count(keep);
    if (threadIdx.x == 0) block_counts[blockIdx.x] = (unsigned int)c;
}

// Pass 2: exclusive scan of the CTA counts in place, continued from *running (rows kept in earlier chunks).
__global__ void __launch_bounds__(1024) ply_scan_kernel(unsigned int* __restrict__ block_counts, unsigned int n_blocks,
                                                        unsigned int* __restrict__ running)
{
    __shared__ unsigned int s_w[32];
    __shared__ unsigned int s_carry;
    const unsigned int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_carry = *running;
    __syncthreads();
    for (unsigned int b0 = 0; b0 < n_blocks; b0 += 1024) {
        const unsigned int i = b0 + tid;
        const unsigned int v = i < n_blocks ? block_counts[i] : 0u;
        unsigned int incl = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const unsigned int t = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= (unsigned)d) incl += t; }
        if (lane == 31) s_w[warp] = incl;
        __syncthreads();
        unsigned int wbase = 0, total = 0;
        for (int k = 0; k < 32; ++k) { const unsigned int t = s_w[k]; if (k < (int)warp) wbase += t; total += t; }
        const unsigned int carry = s_carry;
        if (i < n_blocks) block_counts[i] = carry + wbase + incl - v;
        __syncthreads();
        if (tid == 0) s_carry = carry + total;
        __syncthreads();
    }
    if (tid == 0) *running = s_carry;
}

// Pass 3: decode PLY_ROWS rows per CTA through shared memory and write the kept ones, in file order.
// slot_dst[j] = PointData float index (0..61) the j-th property of a row goes to, or -1 (ignored).
__global__ void __launch_bounds__(PLY_ROWS) ply_decode_kernel(const float* __restrict__ rows, unsigned int n_rows, unsigned int stride,
                                                              const int* __restrict__ slot_dst, float crop,
                                                              const unsigned int* __restrict__ block_offsets,
                                                              float* __restrict__ aos_out, float* __restrict__ sh_out)
{
    extern __shared__ float s_mem[];
    float* s_in = s_mem;                                    // [PLY_ROWS][stride] raw properties
    float* s_pt = s_mem + (size_t)PLY_ROWS * stride;        // [PLY_ROWS][63]     PointData rows (63: odd stride, no bank conflicts)
    __shared__ unsigned int s_warp[PLY_ROWS / 32];
    const unsigned int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned int base = blockIdx.x * PLY_ROWS;
    const unsigned int nrow = min((unsigned int)PLY_ROWS, n_rows - base);
    const float* src = rows + (size_t)base * stride;
    for (unsigned int k = tid; k < nrow * stride; k += PLY_ROWS) s_in[k] = __ldg(src + k);
    __syncthreads();

    bool keep = false;
    float* p = s_pt + tid * 63;
    if (tid < nrow) {
        // PointData defaults (src/ply_loader.h:7-17): opacity 1, scale 0.01, rot (1,0,0,0), everything else 0
        for (int k = 0; k < PLY_POINT_FLOATS; ++k) p[k] = 0.0f;
        p[9] = 1.0f; p[10] = p[11] = p[12] = 0.01f; p[13] = 1.0f;
        const float* r = s_in + tid * stride;
        for (unsigned int j = 0; j < stride; ++j) {
            const int d = __ldg(slot_dst + j);
            if (d < 0) continue;
            float v = r[j];
            if (d == 9) v = __fdiv_rn(1.0f, __fadd_rn(1.0f, expf_libm(-v)));          // sigmoid, src/ply_loader.cpp:116
            else if (d >= 10 && d <= 12) v = expf_libm(v);                              // :117-119
            p[d] = v;
        }
        if (!(p[6] == 0.0f && p[7] == 0.0f && p[8] == 0.0f)) {                          // :133-139 (all-zero DC stays black)
#pragma unroll
            for (int k = 6; k < 9; ++k)
                p[k] = fmaxf(0.0f, fminf(1.0f, __fadd_rn(0.5f, __fmul_rn(0.28209479177387814f, p[k]))));
        }
        keep = inside_crop(p[0], p[1], p[2], crop);
    }
    // order-preserving compaction inside the CTA
    const unsigned int bal = __ballot_sync(0xFFFFFFFFu, keep);
    if (lane == 0) s_warp[warp] = __popc(bal);
    __syncthreads();
    unsigned int before = __popc(bal & ((1u << lane) - 1u));
    for (unsigned int k = 0; k < warp; ++k) before += s_warp[k];
    if (!keep) return;
    const size_t o = (size_t)block_offsets[blockIdx.x] + before;
    float* a = aos_out + o * 14;            // SplatInstance: rotation wxyz, scale, position, colour + opacity
    a[0] = p[13]; a[1] = p[14]; a[2] = p[15]; a[3] = p[16];
    a[4] = p[10]; a[5] = p[11]; a[6] = p[12];
    a[7] = p[0]; a[8] = p[1]; a[9] = p[2];
    a[10] = p[6]; a[11] = p[7]; a[12] = p[8]; a[13] = p[9];
    if (sh_out) {
        float* sh = sh_out + o * 45;
        for (int k = 0; k < 45; ++k) sh[k] = p[17 + k];
    }
}

size_t ply_decode_smem_bytes(unsigned int stride) { return (size_t)PLY_ROWS * (stride + 63) * sizeof(float); }

cudaError_t launch_ply_decode_chunk(const float* rows_dev, unsigned int n_rows, unsigned int stride, const int* slot_dst_dev,
                                    int jx, int jy, int jz, float crop, unsigned int* block_counts, unsigned int* running,
                                    float* aos_out, float* sh_out, cudaStream_t s)
{
    if (n_rows == 0) return cudaSuccess;
    const unsigned int n_blocks = (n_rows + PLY_ROWS - 1) / PLY_ROWS;
    const size_t smem = ply_decode_smem_bytes(stride);
    if (smem > 48 * 1024) {
        const cudaError_t e = cudaFuncSetAttribute(ply_decode_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    ply_count_kernel<<<n_blocks, PLY_ROWS, 0, s>>>(rows_dev, n_rows, stride, jx, jy, jz, crop, block_counts);
    ply_scan_kernel<<<1, 1024, 0, s>>>(block_counts, n_blocks, running);
    ply_decode_kernel<<<n_blocks, PLY_ROWS, smem, s>>>(rows_dev, n_rows, stride, slot_dst_dev, crop, block_counts, aos_out, sh_out);
    return cudaGetLastError();
}
