// C entry points over the host classes, for the Python tests (tests/test_host_cpu.py,
// tests/test_gpu_host.py). Not part of the drop-in surface: that is the C++ classes themselves.
#include <cstring>

#include "instanced_splat_renderer.h"
#include "ply_loader.h"
#include "trackball_camera.h"

extern "C" {

// ---- PLYLoader
void* host_ply_load(const char* path, unsigned* n) {
    auto* v = new std::vector<PointData>();
    if (!PLYLoader::load(path, *v)) { delete v; if (n) *n = 0; return nullptr; }
    if (n) *n = (unsigned)v->size();
    return v;
}
const float* host_ply_data(void* h) { return reinterpret_cast<const float*>(static_cast<std::vector<PointData>*>(h)->data()); }
void host_ply_free(void* h) { delete static_cast<std::vector<PointData>*>(h); }

// ---- TrackballCamera
void* host_cam_create() { return new TrackballCamera(); }
void host_cam_destroy(void* c) { delete static_cast<TrackballCamera*>(c); }
void host_cam_set_viewport(void* c, int w, int h) { static_cast<TrackballCamera*>(c)->setViewportSize(w, h); }
void host_cam_set_position(void* c, float x, float y, float z) { static_cast<TrackballCamera*>(c)->setPosition({x, y, z}); }
void host_cam_set_target(void* c, float x, float y, float z) { static_cast<TrackballCamera*>(c)->setTarget({x, y, z}); }
void host_cam_set_distance(void* c, float d) { static_cast<TrackballCamera*>(c)->setDistance(d); }
void host_cam_mouse_down(void* c, float x, float y, int b) { static_cast<TrackballCamera*>(c)->handleMouseDown(x, y, b); }
void host_cam_mouse_move(void* c, float x, float y) { static_cast<TrackballCamera*>(c)->handleMouseMove(x, y); }
void host_cam_mouse_up(void* c) { static_cast<TrackballCamera*>(c)->handleMouseUp(); }
void host_cam_scroll(void* c, float d) { static_cast<TrackballCamera*>(c)->handleScroll(d); }
void host_cam_view(void* c, float o[16]) { const simd_float4x4 m = static_cast<TrackballCamera*>(c)->getViewMatrix(); std::memcpy(o, &m, 64); }
void host_cam_proj(void* c, float o[16]) { const simd_float4x4 m = static_cast<TrackballCamera*>(c)->getProjectionMatrix(); std::memcpy(o, &m, 64); }
void host_cam_position(void* c, float o[3]) { const simd_float3 p = static_cast<TrackballCamera*>(c)->getPosition(); o[0] = p.x; o[1] = p.y; o[2] = p.z; }
void host_cam_target(void* c, float o[3]) { const simd_float3 p = static_cast<TrackballCamera*>(c)->getTarget(); o[0] = p.x; o[1] = p.y; o[2] = p.z; }

// ---- InstancedSplatRenderer
void* host_renderer_create(const char* ply_path) { return new InstancedSplatRenderer(ply_path); }
void* host_renderer_create_device_decode(const char* ply_path) { return new InstancedSplatRenderer(ply_path, SplatDecode::Device); }
void host_renderer_destroy(void* r) { delete static_cast<InstancedSplatRenderer*>(r); }
int host_renderer_initialize(void* r, int cuda_device) { SplatDevice d{cuda_device}; return static_cast<InstancedSplatRenderer*>(r)->initialize(&d) ? 1 : 0; }
int host_renderer_point_count(void* r) { return static_cast<InstancedSplatRenderer*>(r)->getPointCount(); }
const float* host_renderer_instances(void* r) { return reinterpret_cast<const float*>(static_cast<InstancedSplatRenderer*>(r)->instances().data()); }
void host_renderer_set_options(void* r, const splat_opts* o) { static_cast<InstancedSplatRenderer*>(r)->setOptions(*o); }
void host_renderer_render(void* r, void* stream, void* fb_dev, const float view[16], const float proj[16], float w, float h) {
    simd_float4x4 V, P; std::memcpy(&V, view, 64); std::memcpy(&P, proj, 64);
    static_cast<InstancedSplatRenderer*>(r)->render(stream, fb_dev, V, P, w, h);
}
void* host_renderer_context(void* r) { return static_cast<InstancedSplatRenderer*>(r)->context(); }

}  // extern "C"
