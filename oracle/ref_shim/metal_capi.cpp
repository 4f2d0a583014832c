// Oracle-A: executes the reference's OWN shader statements on the CPU.
// TEST INFRASTRUCTURE ONLY (see oracle/oracle_b.h). Built only where /root/reference exists;
// the .metal files are #included from there and never copied into this repository.
//
// What is the reference's: every statement of vertex_main, fragment_main /
// fragment_accumulate and tile_sort_composite / compute_sort_composite
// (shaders/gaussian_splat_tile.metal, shaders/gaussian_splat_50layer.metal), and of the MLAB
// mode's vertex_main, fragment_main and resolve_main (shaders/gaussian_splat.metal).
// What is ours (the fixed-function stages Metal supplies): clip 0 <= z <= w,
// viewport transform, pixel-centre coverage of the two triangles of each instanced quad,
// barycentric interpolation of the varyings, per-pixel list plumbing,
// instance-order submission (src/instanced_splat_renderer.mm:496-499 draws one instanced call).
#include <cmath>
#include <cstdint>
#include <cstring>
#include <cstdio>
#include <thread>
#include <vector>

#include <metal_stdlib>

namespace ref_tile {
#include "gaussian_splat_tile.metal"
}
#undef MAX_FRAGMENTS_PER_PIXEL
namespace ref_layer {
#include "gaussian_splat_50layer.metal"
}
#undef MAX_FRAGMENTS_PER_PIXEL
namespace ref_mlab {
#include "gaussian_splat.metal"
}
#undef NUM_OIT_LAYERS
#undef constant
#undef device
#undef thread
#undef threadgroup
#undef vertex
#undef fragment
#undef kernel

namespace {

using metal::float2; using metal::float4; using metal::half4; using metal::uint2;

const int K_TILE = 32, K_LAYER = 50;
int g_dbg_x = -1, g_dbg_y = -1;

// simd_mul(proj, view) as in src/instanced_splat_renderer.mm:453 (columns, k left to right)
template <typename M> void load_mat(M& m, const float* p) { for (int c = 0; c < 4; ++c) for (int r = 0; r < 4; ++r) m[c][r] = p[c * 4 + r]; }
template <typename M> void mul_mat(M& o, const M& a, const M& b) {
    for (int j = 0; j < 4; ++j) for (int r = 0; r < 4; ++r)
        o[j][r] = ((a[0][r] * b[j][0] + a[1][r] * b[j][1]) + a[2][r] * b[j][2]) + a[3][r] * b[j][3];
}

struct Pt { double x, y; };
inline double edge(const Pt& a, const Pt& b, double px, double py) { return (b.x - a.x) * (py - a.y) - (b.y - a.y) * (px - a.x); }

// Rasterise the instanced quads in instance order and hand every covered pixel centre to `frag`
// with linearly interpolated varyings (w = 1 on every vertex: affine interpolation is exact).
template <typename VOut, typename VS, typename FS>
void raster(uint32_t n, int w, int h, VS&& vs, FS&& frag) {
    std::vector<uint32_t> stamp((size_t)w * h, 0);
    for (uint32_t i = 0; i < n; ++i) {
        VOut v[6];
        for (int k = 0; k < 6; ++k) v[k] = vs((unsigned)k, i);
        if (v[0].depth < 0.0f) continue;                         // culled in the vertex stage: zero-area quad at the origin
        const float z = v[0].position.z;                         // all six share z and w = 1
        if (!(z >= 0.0f && z <= 1.0f)) continue;                 // Metal clip volume 0 <= z <= w
        Pt p[6];
        for (int k = 0; k < 6; ++k) {
            p[k].x = ((double)v[k].position.x * 0.5 + 0.5) * w;
            p[k].y = (0.5 - 0.5 * (double)v[k].position.y) * h;  // window y is down
        }
        for (int t = 0; t < 2; ++t) {
            const Pt &a = p[3 * t], &b = p[3 * t + 1], &c = p[3 * t + 2];
            const VOut &va = v[3 * t], &vb = v[3 * t + 1], &vc = v[3 * t + 2];
            const double area = edge(a, b, c.x, c.y);
            if (area == 0.0) continue;
            const double xmin = std::fmin(a.x, std::fmin(b.x, c.x)), xmax = std::fmax(a.x, std::fmax(b.x, c.x));
            const double ymin = std::fmin(a.y, std::fmin(b.y, c.y)), ymax = std::fmax(a.y, std::fmax(b.y, c.y));
            int i0 = (int)std::ceil(xmin - 0.5), i1 = (int)std::floor(xmax - 0.5);
            int j0 = (int)std::ceil(ymin - 0.5), j1 = (int)std::floor(ymax - 0.5);
            if (i0 < 0) i0 = 0; if (j0 < 0) j0 = 0; if (i1 > w - 1) i1 = w - 1; if (j1 > h - 1) j1 = h - 1;
            for (int j = j0; j <= j1; ++j) for (int ii = i0; ii <= i1; ++ii) {
                const double px = ii + 0.5, py = j + 0.5;
                double l0 = edge(b, c, px, py) / area, l1 = edge(c, a, px, py) / area, l2 = edge(a, b, px, py) / area;
                if (l0 < 0 || l1 < 0 || l2 < 0) continue;        // closed edges (measure-zero difference from top-left rule)
                uint32_t& st = stamp[(size_t)j * w + ii];
                if (st == i + 1) continue;                       // shared diagonal: one fragment per pixel per quad
                st = i + 1;
                VOut in;
                in.position = float4((float)px, (float)py, z, 1.0f);
                in.color = va.color;                             // identical on all vertices
                in.uv = float2((float)(l0 * va.uv.x + l1 * vb.uv.x + l2 * vc.uv.x),
                               (float)(l0 * va.uv.y + l1 * vb.uv.y + l2 * vc.uv.y));
                in.depth = va.depth;
                if (ii == g_dbg_x && j == g_dbg_y) std::fprintf(stderr, "ora dbg splat %u uv %.9g %.9g depth %.9g\n", i, in.uv.x, in.uv.y, in.depth);
                try { frag(in, (unsigned)ii, (unsigned)j); } catch (const metal::fragment_discarded&) {}
            }
        }
    }
}

void store(const std::vector<half4>& img, uint16_t* fb) {
    for (size_t p = 0; p < img.size(); ++p) {
        const _Float16 c[4] = {img[p].x.v, img[p].y.v, img[p].z.v, img[p].w.v};
        std::memcpy(fb + 4 * p, c, 8);
    }
}

}  // namespace

extern "C" {

// mode 0: gaussian_splat_tile.metal (K = 32), mode 1: gaussian_splat_50layer.metal (K = 50),
// mode 2: gaussian_splat.metal (MLAB, 6-layer K-buffer).
// fb_half4: w*h*4 half bit patterns; counts (nullable): fragments that ARRIVED per pixel
// (can exceed K: the shaders keep counting, gaussian_splat_tile.metal:200-206).
int ora_render(const float* instances, uint32_t n, const float* view16, const float* proj16, int w, int h,
               int mode, uint16_t* fb_half4, uint32_t* counts) {
    static_assert(sizeof(ref_tile::SplatInstance) == 56 && sizeof(ref_layer::SplatInstance) == 56, "SplatInstance layout");
    static_assert(sizeof(ref_layer::PixelFragmentList) == 504, "4 + 50*10 bytes per pixel");
    const size_t P = (size_t)w * h;
    std::vector<half4> img(P);
    if (mode == 0) {
        using namespace ref_tile;
        Uniforms U;
        load_mat(U.viewMatrix, view16); load_mat(U.projectionMatrix, proj16);
        mul_mat(U.viewProjectionMatrix, U.projectionMatrix, U.viewMatrix);
        U.viewportSize = float2((float)w, (float)h);
        const SplatInstance* inst = reinterpret_cast<const SplatInstance*>(instances);
        std::vector<metal::atomic_uint> counter(P); std::memset(counter.data(), 0, P * sizeof(metal::atomic_uint));
        std::vector<FragmentData> buf(P * K_TILE);
        raster<VertexOut>(n, w, h,
            [&](unsigned vid, unsigned iid) { return vertex_main(vid, iid, inst, U); },
            [&](const VertexOut& in, unsigned x, unsigned y) {
                const size_t p = (size_t)y * w + x;                // D1: one list per pixel
                fragment_main(in, &counter[p], &buf[p * K_TILE], uint2(x, y));
            });
        metal::texture2d<metal::half, metal::access::write> tex{img.data(), (unsigned)w, (unsigned)h};
        for (unsigned y = 0; y < (unsigned)h; ++y) for (unsigned x = 0; x < (unsigned)w; ++x) {
            const size_t p = (size_t)y * w + x;
            tile_sort_composite(tex, &buf[p * K_TILE], &counter[p], uint2(x, y), uint2(x / 16, y / 16), uint2(16, 16));
            if (counts) counts[p] = counter[p].v;
        }
    } else if (mode == 1) {
        using namespace ref_layer;
        Uniforms U;
        load_mat(U.viewMatrix, view16); load_mat(U.projectionMatrix, proj16);
        mul_mat(U.viewProjectionMatrix, U.projectionMatrix, U.viewMatrix);
        U.viewportSize = float2((float)w, (float)h);
        const SplatInstance* inst = reinterpret_cast<const SplatInstance*>(instances);
        std::vector<PixelFragmentList> lists(P); std::memset((void*)lists.data(), 0, P * sizeof(PixelFragmentList));
        raster<VertexOut>(n, w, h,
            [&](unsigned vid, unsigned iid) { return vertex_main(vid, iid, inst, U); },
            [&](const VertexOut& in, unsigned, unsigned) { fragment_accumulate(in, lists.data(), U); });
        metal::texture2d<metal::half, metal::access::write> tex{img.data(), (unsigned)w, (unsigned)h};
        for (unsigned y = 0; y < (unsigned)h; ++y) for (unsigned x = 0; x < (unsigned)w; ++x) {
            if (counts) counts[(size_t)y * w + x] = lists[(size_t)y * w + x].count.v;   // before the in-place sort
            compute_sort_composite(lists.data(), tex, U, uint2(x, y));
        }
    } else if (mode == 2) {
        // MLAB: K-buffer pass over 8 attachments cleared to (0,0,0,1), then the resolve pass
        // (src/instanced_splat_renderer.mm:530-577). Raster order groups = primitive (instance) order
        // per pixel, which is the order raster() submits fragments in. The image is resolve_main's
        // return value (the SrcAlpha blend over the drawable is not applied).
        using namespace ref_mlab;
        static_assert(sizeof(ref_mlab::SplatInstance) == 56, "SplatInstance layout");
        Uniforms U;
        load_mat(U.viewMatrix, view16); load_mat(U.projectionMatrix, proj16);
        mul_mat(U.viewProjectionMatrix, U.projectionMatrix, U.viewMatrix);
        U.viewportSize = float2((float)w, (float)h);
        const SplatInstance* inst = reinterpret_cast<const SplatInstance*>(instances);
        const half4 clear(metal::half(0.0f), metal::half(0.0f), metal::half(0.0f), metal::half(1.0f));
        KBuffer init;
        init.layer0 = init.layer1 = init.layer2 = init.layer3 = init.layer4 = init.layer5 = clear;
        init.depths01 = init.depths23 = clear;
        std::vector<KBuffer> kb(P, init);
        std::vector<uint32_t> arrived(counts ? P : 0, 0);
        raster<VertexOut>(n, w, h,
            [&](unsigned vid, unsigned iid) { return vertex_main(vid, iid, inst, U); },
            [&](const VertexOut& in, unsigned x, unsigned y) {
                const size_t p = (size_t)y * w + x;
                kb[p] = fragment_main(in, kb[p]);
                if (counts) ++arrived[p];
            });
        for (size_t p = 0; p < P; ++p) {
            KBufferNoDepth k;
            k.layer0 = kb[p].layer0; k.layer1 = kb[p].layer1; k.layer2 = kb[p].layer2;
            k.layer3 = kb[p].layer3; k.layer4 = kb[p].layer4; k.layer5 = kb[p].layer5;
            img[p] = resolve_main(k).color;
            if (counts) counts[p] = arrived[p];
        }
    } else return -1;
    store(img, fb_half4);
    (void)K_LAYER;
    return 0;
}

void ora_debug_pixel(int x, int y) { g_dbg_x = x; g_dbg_y = y; }

int ora_max_fragments(int mode) { return mode == 0 ? K_TILE : K_LAYER; }

}  // extern "C"
