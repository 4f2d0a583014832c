"""ctypes binding of oracle/_ref (the REFERENCE's own sources compiled where they lie).

TEST INFRASTRUCTURE ONLY. Three libraries, built by oracle/Makefile when /root/reference exists:
  libref_metal.so   Oracle-A: the .metal shader text compiled as C++ + a small software rasteriser
  libref_camera.so  src/trackball_camera.mm over ref_shim/simd/simd.h
  libref_ply.so     src/ply_loader.cpp, unmodified
On the GPU box /root/reference is absent and the prebuilt files (shipped with the snapshot) are used.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_REF = os.path.join(_HERE, "_ref")
_libs = {}


def available(name="metal") -> bool:
    return os.path.exists(os.path.join(_REF, f"libref_{name}.so"))


def _lib(name):
    if name not in _libs:
        _libs[name] = C.CDLL(os.path.join(_REF, f"libref_{name}.so"))
    return _libs[name]


def render(instances, view, proj, w, h, mode=0):
    """Runs the reference shader statements. Returns (image uint16[h,w,4] half bits, arrivals uint32[h,w])."""
    L = _lib("metal")
    inst = np.ascontiguousarray(instances, np.float32).reshape(-1, 14)
    v = np.ascontiguousarray(view, np.float32); p = np.ascontiguousarray(proj, np.float32)
    fb = np.zeros((h, w, 4), np.uint16); cnt = np.zeros((h, w), np.uint32)
    L.ora_render.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    rc = L.ora_render(inst.ctypes.data, inst.shape[0], v.ctypes.data, p.ctypes.data, w, h, mode, fb.ctypes.data, cnt.ctypes.data)
    assert rc == 0
    return fb, cnt


def max_fragments(mode):
    return _lib("metal").ora_max_fragments(mode)


class RefCamera:
    """The reference's TrackballCamera (src/trackball_camera.h:5-64)."""

    def __init__(self):
        L = _lib("camera"); self.L = L
        L.refcam_create.restype = C.c_void_p
        for f in ("destroy", "mouse_up"):
            getattr(L, "refcam_" + f).argtypes = [C.c_void_p]
        L.refcam_set_viewport.argtypes = [C.c_void_p, C.c_int, C.c_int]
        for f in ("set_position", "set_target"):
            getattr(L, "refcam_" + f).argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_float]
        L.refcam_set_distance.argtypes = [C.c_void_p, C.c_float]
        L.refcam_scroll.argtypes = [C.c_void_p, C.c_float]
        L.refcam_mouse_down.argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_int]
        L.refcam_mouse_move.argtypes = [C.c_void_p, C.c_float, C.c_float]
        for f in ("view", "proj", "position", "target"):
            getattr(L, "refcam_" + f).argtypes = [C.c_void_p, C.c_void_p]
        self.h = L.refcam_create()

    def __del__(self):
        if getattr(self, "h", None):
            self.L.refcam_destroy(self.h); self.h = None

    def set_viewport(self, w, h): self.L.refcam_set_viewport(self.h, w, h)
    def set_position(self, x, y, z): self.L.refcam_set_position(self.h, x, y, z)
    def set_target(self, x, y, z): self.L.refcam_set_target(self.h, x, y, z)
    def set_distance(self, d): self.L.refcam_set_distance(self.h, d)
    def scroll(self, d): self.L.refcam_scroll(self.h, d)
    def mouse_down(self, x, y, b): self.L.refcam_mouse_down(self.h, x, y, b)
    def mouse_move(self, x, y): self.L.refcam_mouse_move(self.h, x, y)
    def mouse_up(self): self.L.refcam_mouse_up(self.h)

    def _get(self, name, n):
        o = np.zeros(n, np.float32); getattr(self.L, "refcam_" + name)(self.h, o.ctypes.data); return o

    def view(self): return self._get("view", 16)
    def proj(self): return self._get("proj", 16)
    def position(self): return self._get("position", 3)
    def target(self): return self._get("target", 3)


def load_ply(path):
    """The reference's PLYLoader::load. Returns float32[n, 62] (PointData records) or None on failure."""
    L = _lib("ply")
    L.refply_load.restype = C.c_void_p; L.refply_load.argtypes = [C.c_char_p, C.POINTER(C.c_uint)]
    L.refply_data.restype = C.c_void_p; L.refply_data.argtypes = [C.c_void_p]
    L.refply_free.argtypes = [C.c_void_p]
    n = C.c_uint(0)
    h = L.refply_load(os.fsencode(path), C.byref(n))
    if not h:
        return None
    assert L.refply_record_bytes() == 248
    buf = (C.c_float * (n.value * 62)).from_address(L.refply_data(h))
    out = np.frombuffer(buf, np.float32).reshape(n.value, 62).copy()
    L.refply_free(h)
    return out
