// Presentation step after the path: RGBA16F framebuffer -> BGRA8Unorm, the drawable format of the
// reference's layer (src/metal_renderer.mm:58) that its final blit / resolve pass targets
// (src/instanced_splat_renderer.mm:513-527, 557-576). Pure streaming: 8 B read + 4 B written per pixel.
// unorm8(x) = rint(clamp(x, 0, 1) * 255), NaN -> 0 (round to nearest even, as float -> unorm conversion rounds).
#include "splat_internal.cuh"

__device__ __forceinline__ unsigned int unorm8(unsigned short hbits) {
    const float x = __saturatef(__half2float(__ushort_as_half(hbits)));      // clamps, NaN -> 0
    return __float2uint_rn(__fmul_rn(x, 255.0f));
}

__device__ __forceinline__ unsigned int pack_pixel(uint2 p) {
    const unsigned int r = unorm8((unsigned short)(p.x & 0xFFFFu)), g = unorm8((unsigned short)(p.x >> 16));
    const unsigned int b = unorm8((unsigned short)(p.y & 0xFFFFu)), a = unorm8((unsigned short)(p.y >> 16));
    return b | (g << 8) | (r << 16) | (a << 24);                             // bytes in memory: B G R A
}

__global__ void __launch_bounds__(256) pack_bgra8_kernel(const uint2* __restrict__ fb, unsigned int* __restrict__ out,
                                                         size_t n_pixels)
{
    // two pixels per thread: one 16-byte load, one 8-byte store
    const size_t pair = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t p = pair * 2;
    if (p + 1 < n_pixels) {
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(fb) + pair);
        uint2 o;
        o.x = pack_pixel(make_uint2(v.x, v.y));
        o.y = pack_pixel(make_uint2(v.z, v.w));
        reinterpret_cast<uint2*>(out)[pair] = o;
    } else if (p < n_pixels) {
        out[p] = pack_pixel(__ldg(fb + p));
    }
}

void launch_pack_bgra8(const void* fb_half4, void* bgra8, size_t n_pixels, cudaStream_t s)
{
    if (n_pixels == 0) return;
    const size_t pairs = (n_pixels + 1) / 2;
    pack_bgra8_kernel<<<(unsigned int)((pairs + 255) / 256), 256, 0, s>>>(reinterpret_cast<const uint2*>(fb_half4),
                                                                          reinterpret_cast<unsigned int*>(bgra8), n_pixels);
}
