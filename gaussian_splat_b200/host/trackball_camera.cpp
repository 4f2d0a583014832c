// Orbit camera behind the reference's TrackballCamera interface. State and arithmetic follow
// src/trackball_camera.mm (defaults :5-17, setters :19-37, drag handling :39-99, zoom :101-105,
// matrices :127-163) so that view/projection come out bit-identical to the reference's under the
// same compiler flags; tests/test_host_cpu.py checks that against the reference camera compiled on this box.
#include "trackball_camera.h"

TrackballCamera::TrackballCamera()
    : position{0, 0, 5}, target{0, 0, 0}, up{0, -1, 0}, distance(5.0f), viewportWidth(800), viewportHeight(600),
      drag(Drag::None), lastMousePos{0, 0} {}

void TrackballCamera::setViewportSize(int width, int height) { viewportWidth = width; viewportHeight = height; }
void TrackballCamera::setTarget(simd_float3 t) { target = t; }

void TrackballCamera::setPosition(simd_float3 p) {
    position = p;
    distance = simd_length(position - target);
}

void TrackballCamera::setDistance(float d) {
    distance = simd_clamp(d, minDistance, maxDistance);
    position = target + simd_normalize(position - target) * distance;
}

void TrackballCamera::handleMouseDown(float x, float y, int button) {
    lastMousePos = {x, y};
    if (button == 0) drag = Drag::Rotate;
    else if (button == 1 || button == 2) drag = Drag::Pan;
}

void TrackballCamera::handleMouseUp() { drag = Drag::None; }

void TrackballCamera::handleMouseMove(float x, float y) {
    const simd_float2 now = {x, y};
    const simd_float2 delta = lastMousePos - now;
    if (drag == Drag::Rotate) {
        // horizontal drag turns about world +Y, vertical drag about the camera's right axis
        const float yaw = delta.x * rotateSpeed * 0.01f, pitch = delta.y * rotateSpeed * 0.01f;
        const simd_float3 world_up = {0, 1, 0};
        const simd_float3 right = simd_normalize(simd_cross(simd_normalize(target - position), world_up));
        const simd_quatf q = simd_mul(simd_quaternion(-yaw, world_up), simd_quaternion(-pitch, right));
        position = target + simd_act(q, position - target);
        up = simd_normalize(simd_act(q, up));
    } else if (drag == Drag::Pan) {
        const simd_float3 right = simd_normalize(simd_cross(target - position, up));
        const simd_float3 up_vec = simd_normalize(simd_cross(right, target - position));
        const float px = delta.x * panSpeed * distance / viewportHeight;
        const float py = -delta.y * panSpeed * distance / viewportHeight;
        const simd_float3 shift = right * px + up_vec * py;
        position += shift;
        target += shift;
    }
    lastMousePos = now;
}

void TrackballCamera::handleScroll(float delta) { setDistance(distance * std::pow(0.95f, delta * zoomSpeed)); }

simd_float4x4 TrackballCamera::getViewMatrix() const {
    const simd_float3 f = simd_normalize(target - position);
    const simd_float3 s = simd_normalize(simd_cross(f, up));
    const simd_float3 u = simd_cross(s, f);
    return simd_matrix(simd_make_float4(s.x, u.x, -f.x, 0), simd_make_float4(s.y, u.y, -f.y, 0),
                       simd_make_float4(s.z, u.z, -f.z, 0),
                       simd_make_float4(-simd_dot(s, position), -simd_dot(u, position), simd_dot(f, position), 1));
}

simd_float4x4 TrackballCamera::getProjectionMatrix() const {
    const float aspect = (float)viewportWidth / (float)viewportHeight;
    const float fov_rad = (float)(fov * M_PI / 180.0f);          // double product narrowed once, as the reference's call does
    const float ys = 1.0f / std::tan(fov_rad * 0.5f), xs = ys / aspect;
    const float range = farPlane - nearPlane;
    return simd_matrix(simd_make_float4(xs, 0, 0, 0), simd_make_float4(0, ys, 0, 0),
                       simd_make_float4(0, 0, -(farPlane + nearPlane) / range, -1),
                       simd_make_float4(0, 0, -2.0f * farPlane * nearPlane / range, 0));
}
