"""ctypes binding of libsplat_host.so: the C++ host classes that keep the reference's signatures
(gaussian_splat_b200/host/: PLYLoader, TrackballCamera, InstancedSplatRenderer), reached through
host_capi.cpp. Used by the tests and examples; the drop-in surface itself is the C++ headers."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import capi

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libsplat_host.so")
_LIB = None


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise capi.SplatError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'`")
        capi.lib()                               # libsplat_cuda.so first (dependency, same directory)
        L = C.CDLL(LIB_PATH)
        vp, f = C.c_void_p, C.c_float
        L.host_ply_load.restype = vp; L.host_ply_load.argtypes = [C.c_char_p, C.POINTER(C.c_uint)]
        L.host_ply_data.restype = vp; L.host_ply_data.argtypes = [vp]
        L.host_ply_free.argtypes = [vp]
        L.host_cam_create.restype = vp
        for n in ("destroy", "mouse_up"):
            getattr(L, "host_cam_" + n).argtypes = [vp]
        L.host_cam_set_viewport.argtypes = [vp, C.c_int, C.c_int]
        for n in ("set_position", "set_target"):
            getattr(L, "host_cam_" + n).argtypes = [vp, f, f, f]
        L.host_cam_set_distance.argtypes = [vp, f]; L.host_cam_scroll.argtypes = [vp, f]
        L.host_cam_mouse_down.argtypes = [vp, f, f, C.c_int]; L.host_cam_mouse_move.argtypes = [vp, f, f]
        for n in ("view", "proj", "position", "target"):
            getattr(L, "host_cam_" + n).argtypes = [vp, vp]
        L.host_renderer_create.restype = vp; L.host_renderer_create.argtypes = [C.c_char_p]
        L.host_renderer_create_device_decode.restype = vp; L.host_renderer_create_device_decode.argtypes = [C.c_char_p]
        L.host_renderer_destroy.argtypes = [vp]
        L.host_renderer_initialize.argtypes = [vp, C.c_int]
        L.host_renderer_point_count.argtypes = [vp]
        L.host_renderer_instances.restype = vp; L.host_renderer_instances.argtypes = [vp]
        L.host_renderer_set_options.argtypes = [vp, C.POINTER(capi.SplatOpts)]
        L.host_renderer_render.argtypes = [vp, vp, vp, vp, vp, f, f]
        L.host_renderer_context.restype = vp; L.host_renderer_context.argtypes = [vp]
        _LIB = L
    return _LIB


def load_ply(path):
    """PLYLoader::load -> float32[n, 62] (PointData records) or None."""
    L = lib()
    n = C.c_uint(0)
    h = L.host_ply_load(os.fsencode(path), C.byref(n))
    if not h:
        return None
    buf = (C.c_float * (n.value * 62)).from_address(L.host_ply_data(h))
    out = np.frombuffer(buf, np.float32).reshape(n.value, 62).copy()
    L.host_ply_free(h)
    return out


class TrackballCamera:
    def __init__(self):
        self.L = lib(); self.h = self.L.host_cam_create()

    def __del__(self):
        if getattr(self, "h", None):
            self.L.host_cam_destroy(self.h); self.h = None

    def set_viewport(self, w, h): self.L.host_cam_set_viewport(self.h, w, h)
    def set_position(self, x, y, z): self.L.host_cam_set_position(self.h, x, y, z)
    def set_target(self, x, y, z): self.L.host_cam_set_target(self.h, x, y, z)
    def set_distance(self, d): self.L.host_cam_set_distance(self.h, d)
    def scroll(self, d): self.L.host_cam_scroll(self.h, d)
    def mouse_down(self, x, y, b): self.L.host_cam_mouse_down(self.h, x, y, b)
    def mouse_move(self, x, y): self.L.host_cam_mouse_move(self.h, x, y)
    def mouse_up(self): self.L.host_cam_mouse_up(self.h)

    def _get(self, name, n):
        o = np.zeros(n, np.float32); getattr(self.L, "host_cam_" + name)(self.h, o.ctypes.data); return o

    def view(self): return self._get("view", 16)
    def proj(self): return self._get("proj", 16)
    def position(self): return self._get("position", 3)
    def target(self): return self._get("target", 3)


class InstancedSplatRenderer:
    """The C++ class with the reference's constructor / initialize / render / getPointCount."""

    def __init__(self, ply_path: str, device_decode: bool = False):
        self.L = lib()
        make = self.L.host_renderer_create_device_decode if device_decode else self.L.host_renderer_create
        self.h = make(os.fsencode(ply_path))

    def __del__(self):
        if getattr(self, "h", None):
            self.L.host_renderer_destroy(self.h); self.h = None

    def initialize(self, device: int = 0) -> bool:
        return bool(self.L.host_renderer_initialize(self.h, device))

    def point_count(self) -> int:
        return self.L.host_renderer_point_count(self.h)

    def instances(self) -> np.ndarray:
        n = self.point_count()
        if n == 0:
            return np.zeros((0, 14), np.float32)
        buf = (C.c_float * (n * 14)).from_address(self.L.host_renderer_instances(self.h))
        return np.frombuffer(buf, np.float32).reshape(n, 14).copy()

    def set_options(self, opts): self.L.host_renderer_set_options(self.h, C.byref(opts))

    def render(self, stream_ptr, fb_dev_ptr, view, proj, w, h):
        v = np.ascontiguousarray(view, np.float32); p = np.ascontiguousarray(proj, np.float32)
        self.L.host_renderer_render(self.h, stream_ptr, fb_dev_ptr, v.ctypes.data, p.ctypes.data, float(w), float(h))
