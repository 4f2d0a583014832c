// C entry points over the reference's own TrackballCamera (src/trackball_camera.{h,mm},
// compiled where it lies under /root/reference against ref_shim/simd/simd.h).
// TEST INFRASTRUCTURE ONLY: pins the host-side camera restatement (gaussian_splat_b200/host).
#include "trackball_camera.h"
#include <cstring>

extern "C" {

void* refcam_create() { return new TrackballCamera(); }
void refcam_destroy(void* c) { delete static_cast<TrackballCamera*>(c); }
void refcam_set_viewport(void* c, int w, int h) { static_cast<TrackballCamera*>(c)->setViewportSize(w, h); }
void refcam_set_position(void* c, float x, float y, float z) { static_cast<TrackballCamera*>(c)->setPosition(simd_make_float3(x, y, z)); }
void refcam_set_target(void* c, float x, float y, float z) { static_cast<TrackballCamera*>(c)->setTarget(simd_make_float3(x, y, z)); }
void refcam_set_distance(void* c, float d) { static_cast<TrackballCamera*>(c)->setDistance(d); }
void refcam_mouse_down(void* c, float x, float y, int button) { static_cast<TrackballCamera*>(c)->handleMouseDown(x, y, button); }
void refcam_mouse_move(void* c, float x, float y) { static_cast<TrackballCamera*>(c)->handleMouseMove(x, y); }
void refcam_mouse_up(void* c) { static_cast<TrackballCamera*>(c)->handleMouseUp(); }
void refcam_scroll(void* c, float d) { static_cast<TrackballCamera*>(c)->handleScroll(d); }
void refcam_view(void* c, float out16[16]) { const simd_float4x4 m = static_cast<TrackballCamera*>(c)->getViewMatrix(); std::memcpy(out16, &m, 64); }
void refcam_proj(void* c, float out16[16]) { const simd_float4x4 m = static_cast<TrackballCamera*>(c)->getProjectionMatrix(); std::memcpy(out16, &m, 64); }
void refcam_position(void* c, float out3[3]) { const simd_float3 p = static_cast<TrackballCamera*>(c)->getPosition(); out3[0] = p.x; out3[1] = p.y; out3[2] = p.z; }
void refcam_target(void* c, float out3[3]) { const simd_float3 p = static_cast<TrackballCamera*>(c)->getTarget(); out3[0] = p.x; out3[1] = p.y; out3[2] = p.z; }

}  // extern "C"
