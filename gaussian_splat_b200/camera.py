"""View / projection matrices as the reference's TrackballCamera produces them
(src/trackball_camera.mm:127-163: GL-style lookAt and perspective, column-major float[16]).

All arithmetic is done in float32, one rounding per operation, in the reference's order.
"""
from __future__ import annotations

import math

import numpy as np

f32 = np.float32


def _norm(v):
    l = np.sqrt(f32(f32(f32(v[0] * v[0]) + f32(v[1] * v[1])) + f32(v[2] * v[2])))
    return np.array([v[0] / l, v[1] / l, v[2] / l], f32)


def _cross(a, b):
    return np.array([f32(a[1] * b[2]) - f32(a[2] * b[1]), f32(a[2] * b[0]) - f32(a[0] * b[2]),
                     f32(a[0] * b[1]) - f32(a[1] * b[0])], f32)


def _dot(a, b):
    return f32(f32(f32(a[0] * b[0]) + f32(a[1] * b[1])) + f32(a[2] * b[2]))


def look_at(eye, center, up) -> np.ndarray:
    """makeLookAt (src/trackball_camera.mm:136-148)."""
    eye, center, up = (np.asarray(v, f32) for v in (eye, center, up))
    f = _norm(center - eye)
    s = _norm(_cross(f, up))
    u = _cross(s, f)
    o = np.zeros(16, f32)
    o[0], o[1], o[2] = s[0], u[0], -f[0]
    o[4], o[5], o[6] = s[1], u[1], -f[1]
    o[8], o[9], o[10] = s[2], u[2], -f[2]
    o[12], o[13], o[14], o[15] = -_dot(s, eye), -_dot(u, eye), _dot(f, eye), f32(1)
    return o


def perspective(fov_deg, w, h, z_near=0.1, z_far=1000.0) -> np.ndarray:
    """makePerspective (src/trackball_camera.mm:131-134,150-163)."""
    aspect = f32(w) / f32(h)
    fov = f32(float(f32(fov_deg)) * math.pi / float(f32(180.0)))
    y_scale = f32(1) / f32(math.tan(float(fov * f32(0.5))))
    x_scale = y_scale / aspect
    zn, zf = f32(z_near), f32(z_far)
    z_range = zf - zn
    o = np.zeros(16, f32)
    o[0], o[5] = x_scale, y_scale
    o[10] = -(zf + zn) / z_range
    o[11] = f32(-1)
    o[14] = f32(f32(f32(-2) * zf) * zn) / z_range
    return o


def app_camera(w, h, eye=(0.0, 2.0, 5.0), target=(0.0, 0.0, 0.0), up=(0.0, -1.0, 0.0), fov=45.0):
    """The app's camera: position (0,2,5), target origin, default up (0,-1,0), fov 45, near 0.1,
    far 1000 (src/main.mm:55-58; src/trackball_camera.mm:5-17)."""
    return look_at(eye, target, up), perspective(fov, w, h)
