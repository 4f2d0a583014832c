// Host-side mirror of the reference camera interface (src/trackball_camera.h:5-64): same class,
// method names, public tunables and defaults, over simd_compat.h instead of Apple's <simd/simd.h>.
#pragma once
#include "simd_compat.h"

class TrackballCamera {
public:
    TrackballCamera();

    void setViewportSize(int width, int height);
    void setTarget(simd_float3 target);
    void setPosition(simd_float3 position);
    void setDistance(float distance);

    // button 0 = rotate, 1 or 2 = pan (src/trackball_camera.mm:39-47)
    void handleMouseDown(float x, float y, int button);
    void handleMouseMove(float x, float y);
    void handleMouseUp();
    void handleScroll(float delta);

    simd_float4x4 getViewMatrix() const;        // GL-style lookAt, column-major
    simd_float4x4 getProjectionMatrix() const;  // GL-style perspective: NDC z in [-1, 1]
    simd_float3 getPosition() const { return position; }
    simd_float3 getTarget() const { return target; }

    float rotateSpeed = 1.0f;
    float zoomSpeed = 1.2f;
    float panSpeed = 0.3f;
    float minDistance = 0.1f;
    float maxDistance = 100.0f;

    float fov = 45.0f;
    float nearPlane = 0.1f;
    float farPlane = 1000.0f;

private:
    simd_float3 position, target, up;
    float distance;
    int viewportWidth, viewportHeight;
    enum class Drag { None, Rotate, Pan } drag;
    simd_float2 lastMousePos;
};
