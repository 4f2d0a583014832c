// Host-side mirror of the reference renderer interface (src/instanced_splat_renderer.h:6-34): same
// struct, class, method names and argument order. The three void* keep their roles:
//   initialize(device)        MTLDevice*            -> const SplatDevice* (nullptr = CUDA device 0)
//   render(commandBuffer,     MTLCommandBuffer*     -> cudaStream_t (nullptr = default stream)
//          drawableTexture,   CAMetalDrawable tex   -> device pointer to w*h half4 (8 B / pixel)
//          ...)
// render() only ENQUEUES on the stream and returns, as the reference only encodes
// (src/instanced_splat_renderer.mm:424-528; src/metal_renderer.mm:131 commits later).
#pragma once
#include <string>
#include <vector>

#include "../../include/splat_cuda.h"
#include "simd_compat.h"

struct SplatInstance {
    float rotation[4];  // w x y z
    float scale[3];
    float position[3];
    float color[4];     // r g b opacity
};
static_assert(sizeof(SplatInstance) == 56, "SplatInstance is 56 packed bytes (src/instanced_splat_renderer.h:6-11)");

struct SplatDevice { int cuda_device; };
// Where the PLY vertex table is decoded (activations, DC -> colour, crop): on the host as the reference
// does (src/ply_loader.cpp:91-146 + src/instanced_splat_renderer.mm:360-388), or by the device kernel behind
// splat_upload_ply_rows during initialize() (same results, bit for bit).
enum class SplatDecode { Host, Device };

class InstancedSplatRenderer {
public:
    InstancedSplatRenderer(std::string filepath);
    InstancedSplatRenderer(std::string filepath, SplatDecode where);     // addition: Device defers the decode to initialize()
    ~InstancedSplatRenderer();

    bool initialize(void* device);

    void render(void* commandBuffer, void* drawableTexture, const simd_float4x4& viewMatrix,
                const simd_float4x4& projectionMatrix, float viewportWidth, float viewportHeight);

    int getPointCount() const { return (int)_count; }

    // --- additions (no counterpart in the reference) ---
    // mode / K / depth order / SH degree / tile-row band. Defaults = what the reference app runs: LAYER50, K 50, the
    // shader's literal (descending) depth order (src/instanced_splat_renderer.mm:38; gaussian_splat_50layer.metal:201).
    void setOptions(const splat_opts& opts);
    // Per-stage ms of the last frame rendered with SPLAT_FLAG_TIMING (synchronises).
    bool stageTimes(float ms[8]);
    // Host copy of the instances; with SplatDecode::Device it is read back from the device on first use.
    const std::vector<SplatInstance>& instances();
    splat_ctx* context();

private:
    std::vector<SplatInstance> _instances;
    size_t _count = 0;
    void decodeOnHost(const std::string& filepath);
    InstancedSplatRenderer(const InstancedSplatRenderer&) = delete;
    InstancedSplatRenderer& operator=(const InstancedSplatRenderer&) = delete;
    class Impl;
    Impl* impl;
};
