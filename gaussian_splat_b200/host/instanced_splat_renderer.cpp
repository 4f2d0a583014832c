// InstancedSplatRenderer over libsplat_cuda (include/splat_cuda.h). Constructor, crop and upload
// follow src/instanced_splat_renderer.mm:339-393 (PLY -> SplatInstance, |x|,|y|,|z| < 5 crop) and
// :399-422 (one upload at initialize); render follows :424-528 (silent return on missing handles or
// an empty scene, uniforms from the caller's matrices, whole drawable overwritten).
#include "instanced_splat_renderer.h"

#include <cmath>
#include <iostream>

#include "ply_loader.h"

class InstancedSplatRenderer::Impl {
public:
    splat_ctx* ctx = nullptr;
    // What an unmodified caller gets is what the reference app runs: the 50-layer mode (use50LayerMode = true,
    // src/instanced_splat_renderer.mm:38) with the shader's literal, descending depth order
    // (shaders/gaussian_splat_50layer.metal:201).
    splat_opts opts{SPLAT_MODE_LAYER50, 0, SPLAT_ORDER_REFERENCE, 0, 0, 0, 0, 0};
    std::vector<float> sh_rest;   // [n][45], kept only so a later setOptions(sh_degree > 0) can upload it
    bool sh_uploaded = false;
    bool device_decode = false;
    PLYLoader::Rows raw;          // undecoded vertex table (SplatDecode::Device), released after initialize()
};

InstancedSplatRenderer::InstancedSplatRenderer(std::string filepath, SplatDecode where) : impl(new Impl()) {
    if (where == SplatDecode::Host) { decodeOnHost(filepath); return; }
    impl->device_decode = true;
    if (!PLYLoader::loadRows(filepath, impl->raw))
        std::cerr << "Failed to load Gaussian splat from: " << filepath << std::endl;
}

InstancedSplatRenderer::InstancedSplatRenderer(std::string filepath) : impl(new Impl()) { decodeOnHost(filepath); }

void InstancedSplatRenderer::decodeOnHost(const std::string& filepath) {
    std::vector<PointData> points;
    if (!PLYLoader::load(filepath, points)) {
        std::cerr << "Failed to load Gaussian splat from: " << filepath << std::endl;
        return;                                        // empty renderer, as the reference leaves it
    }
    _instances.reserve(points.size());
    impl->sh_rest.reserve(points.size() * 45);
    const float r = 5.0f;
    for (const PointData& p : points) {
        if (!(std::fabs(p.x) < r && std::fabs(p.y) < r && std::fabs(p.z) < r)) continue;
        SplatInstance s;
        s.rotation[0] = p.rot_0; s.rotation[1] = p.rot_1; s.rotation[2] = p.rot_2; s.rotation[3] = p.rot_3;
        s.scale[0] = p.scale_x; s.scale[1] = p.scale_y; s.scale[2] = p.scale_z;
        s.position[0] = p.x; s.position[1] = p.y; s.position[2] = p.z;
        s.color[0] = p.r; s.color[1] = p.g; s.color[2] = p.b; s.color[3] = p.opacity;
        _instances.push_back(s);
        impl->sh_rest.insert(impl->sh_rest.end(), p.sh_rest, p.sh_rest + 45);
    }
    _count = _instances.size();
}

InstancedSplatRenderer::~InstancedSplatRenderer() {
    if (impl->ctx) splat_destroy(impl->ctx);
    delete impl;
}

bool InstancedSplatRenderer::initialize(void* device) {
    const int dev = device ? static_cast<const SplatDevice*>(device)->cuda_device : 0;
    if (impl->ctx) { splat_destroy(impl->ctx); impl->ctx = nullptr; }
    if (splat_create(dev, &impl->ctx) != SPLAT_OK) {
        std::cerr << "InstancedSplatRenderer: " << splat_last_error_string() << std::endl;
        return false;
    }
    if (impl->device_decode) {
        uint32_t kept = 0;
        const float crop = 5.0f;                           // src/instanced_splat_renderer.mm:360
        if (impl->raw.n && splat_upload_ply_rows(impl->ctx, impl->raw.rows.data(), impl->raw.n, impl->raw.stride,
                                                 impl->raw.slot_dst.data(), crop, 3, &kept) != SPLAT_OK) {
            std::cerr << "InstancedSplatRenderer: " << splat_last_error_string() << std::endl;
            return false;
        }
        _count = kept;
        _instances.clear();
        impl->sh_uploaded = true;                          // f_rest went up with the decode
        return true;
    }
    if (!_instances.empty() && splat_upload(impl->ctx, _instances.data(), (uint32_t)_instances.size()) != SPLAT_OK) {
        std::cerr << "InstancedSplatRenderer: " << splat_last_error_string() << std::endl;
        return false;
    }
    impl->sh_uploaded = false;
    return true;
}

void InstancedSplatRenderer::setOptions(const splat_opts& o) { impl->opts = o; }
splat_ctx* InstancedSplatRenderer::context() { return impl->ctx; }

const std::vector<SplatInstance>& InstancedSplatRenderer::instances() {
    if (_instances.empty() && _count && impl->device_decode && impl->ctx) {
        _instances.resize(_count);
        size_t bytes = _count * sizeof(SplatInstance);
        if (splat_debug_fetch(impl->ctx, SPLAT_FETCH_INSTANCES, _instances.data(), &bytes) != SPLAT_OK) _instances.clear();
    }
    return _instances;
}

bool InstancedSplatRenderer::stageTimes(float ms[8]) { return impl->ctx && splat_stage_times(impl->ctx, ms) == SPLAT_OK; }

void InstancedSplatRenderer::render(void* commandBuffer, void* drawableTexture, const simd_float4x4& view,
                                    const simd_float4x4& proj, float viewportWidth, float viewportHeight) {
    if (!impl->ctx || !drawableTexture || _count == 0) return;   // src/instanced_splat_renderer.mm:434-440
    if (impl->opts.sh_degree > 0 && !impl->sh_uploaded) {
        if (splat_upload_sh(impl->ctx, impl->sh_rest.data(), (uint32_t)_instances.size(), 3) != SPLAT_OK) {
            std::cerr << "InstancedSplatRenderer: " << splat_last_error_string() << std::endl;
            return;
        }
        impl->sh_uploaded = true;
    }
    int rc = splat_render(impl->ctx, commandBuffer, drawableTexture, &view.columns[0].x, &proj.columns[0].x,
                          (int)viewportWidth, (int)viewportHeight, &impl->opts);
    // An EARLIER frame overflowed the pair buffers (it was rendered empty): the library has grown them and says so once.
    // The reference never drops splats (src/instanced_splat_renderer.mm:496-499), so this frame is simply enqueued again.
    if (rc == SPLAT_ERR_OVERFLOW)
        rc = splat_render(impl->ctx, commandBuffer, drawableTexture, &view.columns[0].x, &proj.columns[0].x,
                          (int)viewportWidth, (int)viewportHeight, &impl->opts);
    if (rc != SPLAT_OK) std::cerr << "InstancedSplatRenderer::render: " << splat_last_error_string() << std::endl;
}
