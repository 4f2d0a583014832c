// Micro-benchmark: how fast can a K1-shaped kernel stream on this B200?  (tuning aid, not product)
// Reads two float4 planes (+ optional third), writes one u32 per element; ITEMS elements per thread.
#include <cstdio>
#include <cuda_runtime.h>
template <int ITEMS, int NPL>
__global__ void __launch_bounds__(256) k(const float4* __restrict__ a, const float4* __restrict__ b, const float4* __restrict__ c,
                                         unsigned* __restrict__ out, unsigned n) {
    const unsigned i0 = blockIdx.x * (256u * ITEMS) + threadIdx.x;
    float4 x[ITEMS], y[ITEMS], z[ITEMS];
#pragma unroll
    for (int k = 0; k < ITEMS; ++k) { const unsigned i = i0 + k * 256u; if (i < n) { x[k] = __ldg(a + i); y[k] = __ldg(b + i); if (NPL > 2) z[k] = __ldg(c + i); } }
#pragma unroll
    for (int k = 0; k < ITEMS; ++k) { const unsigned i = i0 + k * 256u; if (i < n) { float s = x[k].x * y[k].y + x[k].z + y[k].w; if (NPL > 2) s += z[k].x; out[i] = s > 1e30f ? 1u : 0u; } }
}
template <int ITEMS, int NPL> float run(const float4* a, const float4* b, const float4* c, unsigned* o, unsigned n) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const unsigned per = 256u * ITEMS, grid = (n + per - 1) / per;
    for (int w = 0; w < 3; ++w) k<ITEMS, NPL><<<grid, 256>>>(a, b, c, o, n);
    cudaEventRecord(e0);
    for (int w = 0; w < 10; ++w) k<ITEMS, NPL><<<grid, 256>>>(a, b, c, o, n);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); return ms / 10;
}
int main() {
    const unsigned n = 6000000;
    float4 *a, *b, *c; unsigned* o;
    cudaMalloc(&a, n * 16ull); cudaMalloc(&b, n * 16ull); cudaMalloc(&c, n * 16ull); cudaMalloc(&o, n * 4ull);
    cudaMemset(a, 0, n * 16ull); cudaMemset(b, 0, n * 16ull); cudaMemset(c, 0, n * 16ull);
    // also a big buffer to flush L2 is unnecessary: 3 planes = 288 MB > 126 MB L2
#define R(I, P) { float ms = run<I, P>(a, b, c, o, n); printf("items=%d planes=%d  %.1f us  %.0f GB/s\n", I, P, ms * 1e3, (n * (16.0 * P + 4)) / ms / 1e6); }
    R(1, 2) R(2, 2) R(4, 2) R(8, 2) R(1, 3) R(2, 3) R(4, 3)
    return 0;
}
