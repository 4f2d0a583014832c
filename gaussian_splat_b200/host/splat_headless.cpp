// Headless driver: the role of the reference's src/main.mm (construct + initialize once :71-72,
// render once per frame :192-198) and of MetalRenderer (owns the stream, the drawable and the
// timing), without a window. Renders an orbit of frames of a 3DGS PLY and optionally writes the
// last one as a binary PPM.
//
//   splat_headless scene.ply [--width 1920] [--height 1080] [--frames 64] [--mode tile|layer50|mlab]
//                  [--k N] [--order near|reference] [--sh 0..3] [--device 0] [--out frame.ppm] [--decode host|device]
//                  [--trace events.txt]   replay an interaction trace through the reference's camera handlers
//                                         (src/trackball_camera.mm:39-105, driven as src/main.mm:104-141 drives them): lines
//                                         `down x y button` | `move x y` | `up` | `scroll delta` | `frame`
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "instanced_splat_renderer.h"
#include "trackball_camera.h"

// P6 from the BGRA8 drawable image produced by splat_pack_bgra8.
static bool write_ppm(const char* path, const std::vector<unsigned char>& bgra, int w, int h) {
    FILE* f = std::fopen(path, "wb");
    if (!f) return false;
    std::fprintf(f, "P6\n%d %d\n255\n", w, h);
    std::vector<unsigned char> row((size_t)w * 3);
    for (int j = 0; j < h; ++j) {
        for (int i = 0; i < w; ++i) {
            const unsigned char* p = &bgra[((size_t)j * w + i) * 4];
            row[(size_t)i * 3 + 0] = p[2]; row[(size_t)i * 3 + 1] = p[1]; row[(size_t)i * 3 + 2] = p[0];
        }
        std::fwrite(row.data(), 1, row.size(), f);
    }
    std::fclose(f);
    return true;
}

int main(int argc, char** argv) {
    if (argc < 2) { std::fprintf(stderr, "usage: %s scene.ply [options]\n", argv[0]); return 2; }
    int w = 1920, h = 1080, frames = 64, device = 0;
    const char* out = nullptr;
    const char* trace = nullptr;
    SplatDecode decode = SplatDecode::Host;
    splat_opts opts{};                     // the driver's own default stays TILE / near-first (BASELINE config 2); --mode / --order choose
    opts.flags = SPLAT_FLAG_TIMING;
    for (int a = 2; a + 1 < argc; a += 2) {
        const std::string k = argv[a]; const char* v = argv[a + 1];
        if (k == "--width") w = std::atoi(v);
        else if (k == "--height") h = std::atoi(v);
        else if (k == "--frames") frames = std::atoi(v);
        else if (k == "--device") device = std::atoi(v);
        else if (k == "--decode") decode = std::strcmp(v, "device") == 0 ? SplatDecode::Device : SplatDecode::Host;
        else if (k == "--k") opts.max_fragments = std::atoi(v);
        else if (k == "--sh") opts.sh_degree = std::atoi(v);
        else if (k == "--out") out = v;
        else if (k == "--trace") trace = v;
        else if (k == "--mode") opts.mode = std::strcmp(v, "layer50") == 0 ? SPLAT_MODE_LAYER50 : std::strcmp(v, "mlab") == 0 ? SPLAT_MODE_MLAB : SPLAT_MODE_TILE;
        else if (k == "--order") opts.order = std::strcmp(v, "reference") == 0 ? SPLAT_ORDER_REFERENCE : SPLAT_ORDER_NEAR_FIRST;
        else { std::fprintf(stderr, "unknown option %s\n", k.c_str()); return 2; }
    }

    InstancedSplatRenderer renderer(argv[1], decode);
    SplatDevice dev{device};
    if (!renderer.initialize(&dev)) return 1;
    renderer.setOptions(opts);
    std::printf("points: %d\n", renderer.getPointCount());

    TrackballCamera camera;                       // src/main.mm:55-58
    camera.setViewportSize(w, h);
    camera.setTarget(simd_make_float3(0, 0, 0));

    cudaStream_t stream; cudaStreamCreate(&stream);
    void* fb = nullptr;
    if (cudaMalloc(&fb, (size_t)w * h * 8) != cudaSuccess) { std::fprintf(stderr, "cudaMalloc failed\n"); return 1; }
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    double total_ms = 0;
    auto render_frame = [&](bool count) {
        for (int attempt = 0; attempt < 4; ++attempt) {
            cudaEventRecord(e0, stream);
            renderer.render(stream, fb, camera.getViewMatrix(), camera.getProjectionMatrix(), (float)w, (float)h);
            cudaEventRecord(e1, stream);
            cudaEventSynchronize(e1);
            splat_status st;
            if (splat_frame_status(renderer.context(), &st) == SPLAT_OK && st.overflow) continue;   // buffers grown: redo
            float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
            if (count) total_ms += ms;
            return;
        }
    };
    if (trace) {
        // the app's start pose (src/main.mm:55-58), then the recorded events; every `frame` line renders one
        camera.setPosition(simd_make_float3(0, 2, 5));
        FILE* tf = std::fopen(trace, "r");
        if (!tf) { std::fprintf(stderr, "cannot open %s\n", trace); return 1; }
        char word[32]; frames = 0;
        while (std::fscanf(tf, "%31s", word) == 1) {
            float a = 0, b = 0; int btn = 0;
            if (!std::strcmp(word, "down") && std::fscanf(tf, "%f %f %d", &a, &b, &btn) == 3) camera.handleMouseDown(a, b, btn);
            else if (!std::strcmp(word, "move") && std::fscanf(tf, "%f %f", &a, &b) == 2) camera.handleMouseMove(a, b);
            else if (!std::strcmp(word, "up")) camera.handleMouseUp();
            else if (!std::strcmp(word, "scroll") && std::fscanf(tf, "%f", &a) == 1) camera.handleScroll(a);
            else if (!std::strcmp(word, "frame")) { render_frame(frames > 0); ++frames; }
            else { std::fprintf(stderr, "bad trace line at '%s'\n", word); return 2; }
        }
        std::fclose(tf);
        const simd_float3 p = camera.getPosition();
        std::printf("trace: %d frames, camera at %.9g %.9g %.9g\n", frames, p.x, p.y, p.z);
    } else {
        for (int f = 0; f < frames; ++f) {
            const float t = 6.28318530718f * (float)f / (float)std::max(frames, 1);
            camera.setPosition(simd_make_float3(5.0f * std::sin(t), 2.0f, 5.0f * std::cos(t)));
            render_frame(f > 0 || frames == 1);
        }
    }
    const int timed = frames > 1 ? frames - 1 : frames;
    if (timed > 0) std::printf("%d frames, %.3f ms/frame, %.1f frames/s\n", timed, total_ms / timed, 1e3 * timed / total_ms);
    if (out) {
        // the reference presents a BGRA8Unorm drawable (src/metal_renderer.mm:58): pack on the device, copy 4 B/pixel
        void* bgra = nullptr;
        std::vector<unsigned char> host((size_t)w * h * 4);
        if (cudaMalloc(&bgra, host.size()) != cudaSuccess ||
            splat_pack_bgra8(renderer.context(), stream, fb, bgra, w, h) != SPLAT_OK ||
            cudaMemcpyAsync(host.data(), bgra, host.size(), cudaMemcpyDeviceToHost, stream) != cudaSuccess ||
            cudaStreamSynchronize(stream) != cudaSuccess) { std::fprintf(stderr, "BGRA8 pack failed\n"); return 1; }
        cudaFree(bgra);
        if (!write_ppm(out, host, w, h)) { std::fprintf(stderr, "cannot write %s\n", out); return 1; }
        std::printf("wrote %s\n", out);
    }
    cudaFree(fb);
    return 0;
}
