"""ctypes binding of Oracle-B (oracle/liboracle_b.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference leg, never by gaussian_splat_b200.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

MODE_TILE, MODE_LAYER50, MODE_MLAB = 0, 1, 2
ORDER_NEAR_FIRST, ORDER_REFERENCE = 0, 1

RECORD_DTYPE = np.dtype([
    ("cx", "<f4"), ("cy", "<f4"), ("ux", "<f4"), ("uy", "<f4"), ("vx", "<f4"), ("vy", "<f4"),
    ("opacity", "<f4"), ("rg", "<u4"), ("bd", "<u4"), ("z_front", "<f4"),
    ("rect_lo", "<u4"), ("rect_hi", "<u4")])
assert RECORD_DTYPE.itemsize == 48


class Opts(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "mode", "max_fragments", "order", "sh_degree", "tile_row_begin", "tile_row_end",
        "cap_policy", "tile_test")]


def build(force: bool = False) -> str:
    so = os.path.join(_HERE, "liboracle_b.so")
    src = os.path.join(_HERE, "oracle_b.cpp")
    if force or not os.path.exists(so) or (
            os.path.exists(src) and os.path.getmtime(so) < os.path.getmtime(src)):
        subprocess.check_call(["make", "-C", _HERE, "liboracle_b.so"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(build())
        fp, u16p, vp = C.POINTER(C.c_float), C.POINTER(C.c_uint16), C.c_void_p
        L.orb_frame_create.restype = vp
        L.orb_frame_create.argtypes = [vp, C.c_uint32, vp, vp, vp, C.c_int, C.c_int, C.POINTER(Opts)]
        L.orb_frame_destroy.argtypes = [vp]
        L.orb_n.restype = C.c_uint32; L.orb_n.argtypes = [vp]
        L.orb_num_pairs.restype = C.c_uint64; L.orb_num_pairs.argtypes = [vp]
        L.orb_num_tiles.restype = C.c_uint32; L.orb_num_tiles.argtypes = [vp]
        for name in ("records", "tiles_touched", "offsets", "keys_unsorted", "vals_unsorted",
                     "keys_sorted", "vals_sorted", "tile_ranges"):
            f = getattr(L, "orb_" + name); f.restype = vp; f.argtypes = [vp]
        L.orb_blend.restype = C.c_uint64; L.orb_blend.argtypes = [vp, vp, C.c_int]
        L.orb_blend_rows.restype = C.c_uint64; L.orb_blend_rows.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int]
        L.orb_look_at.argtypes = [fp, fp, fp, fp]
        L.orb_perspective.argtypes = [C.c_float, C.c_int, C.c_int, C.c_float, C.c_float, fp]
        L.orb_mat4_mul.argtypes = [fp, fp, fp]
        L.orb_exp_neg_half.restype = C.c_float; L.orb_exp_neg_half.argtypes = [C.c_float]
        L.orb_float_to_half.restype = C.c_uint16; L.orb_float_to_half.argtypes = [C.c_float]
        L.orb_half_to_float.restype = C.c_float; L.orb_half_to_float.argtypes = [C.c_uint16]
        L.orb_sort_pairs.argtypes = [vp, vp, C.c_uint64]
        L.orb_set_threads.argtypes = [C.c_int]
        _LIB = L
    return _LIB


def set_threads(n: int):
    lib().orb_set_threads(int(n))


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def look_at(eye, center, up):
    o = np.zeros(16, np.float32)
    e, c, u = (np.asarray(v, np.float32) for v in (eye, center, up))
    lib().orb_look_at(_fp(e), _fp(c), _fp(u), _fp(o))
    return o


def perspective(fov_deg, w, h, z_near=0.1, z_far=1000.0):
    o = np.zeros(16, np.float32)
    lib().orb_perspective(fov_deg, w, h, z_near, z_far, _fp(o))
    return o


def app_camera(w, h, eye=(0.0, 2.0, 5.0), target=(0.0, 0.0, 0.0), up=(0.0, -1.0, 0.0), fov=45.0):
    """The app's camera (src/main.mm:55-58; src/trackball_camera.mm:5-17)."""
    return look_at(eye, target, up), perspective(fov, w, h)


def _view(ptr, dtype, count):
    if count == 0:
        return np.zeros(0, dtype)
    buf = (C.c_char * (count * np.dtype(dtype).itemsize)).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype, count=count).copy()


class Frame:
    """All stages up to tile ranges for one (scene, camera, options)."""

    def __init__(self, instances, view, proj, w, h, mode=MODE_TILE, max_fragments=None,
                 order=ORDER_NEAR_FIRST, tile_row_begin=0, tile_row_end=0, sh_rest=None, sh_degree=0, tile_test=1):
        L = lib()
        inst = np.ascontiguousarray(instances, np.float32).reshape(-1, 14)
        if max_fragments is None:
            max_fragments = 50 if mode == MODE_LAYER50 else 32
        self.opts = Opts(mode, max_fragments, order, sh_degree, tile_row_begin, tile_row_end, 0, tile_test)
        self.w, self.h = w, h
        v = np.ascontiguousarray(view, np.float32); p = np.ascontiguousarray(proj, np.float32)
        shp = None
        if sh_rest is not None:
            self._sh = np.ascontiguousarray(sh_rest, np.float32); shp = self._sh.ctypes.data
        self._h = L.orb_frame_create(inst.ctypes.data, inst.shape[0], shp, v.ctypes.data, p.ctypes.data,
                                     w, h, C.byref(self.opts))
        self.n = inst.shape[0]
        self.D = int(L.orb_num_pairs(self._h))
        self.T = int(L.orb_num_tiles(self._h))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orb_frame_destroy(self._h); self._h = None

    def _get(self, name, dtype, count):
        return _view(getattr(lib(), "orb_" + name)(self._h), dtype, count)

    @property
    def records(self): return self._get("records", RECORD_DTYPE, self.n)
    @property
    def tiles_touched(self): return self._get("tiles_touched", np.uint32, self.n)
    @property
    def offsets(self): return self._get("offsets", np.uint32, self.n)
    @property
    def keys_unsorted(self): return self._get("keys_unsorted", np.uint64, self.D)
    @property
    def vals_unsorted(self): return self._get("vals_unsorted", np.uint32, self.D)
    @property
    def keys_sorted(self): return self._get("keys_sorted", np.uint64, self.D)
    @property
    def vals_sorted(self): return self._get("vals_sorted", np.uint32, self.D)
    @property
    def tile_ranges(self): return self._get("tile_ranges", np.uint32, self.T * 2).reshape(-1, 2)

    def blend(self, n_threads=1, fb=None, rows=None):
        """Returns (image uint16 [h,w,4] of half bits, number of (pixel,splat) tests). rows = (begin, end): only those
        tile rows (of the frame's band) are blended, the rest of fb is left as it is."""
        if fb is None:
            fb = np.zeros((self.h, self.w, 4), np.uint16)
        if rows is None:
            tests = lib().orb_blend(self._h, fb.ctypes.data, n_threads)
        else:
            tests = lib().orb_blend_rows(self._h, fb.ctypes.data, int(rows[0]), int(rows[1]), n_threads)
        return fb, int(tests)


def half_bits_to_float(img_u16):
    return img_u16.view(np.float16).astype(np.float32)


def sort_pairs(keys, vals):
    k = np.ascontiguousarray(keys, np.uint64).copy(); v = np.ascontiguousarray(vals, np.uint32).copy()
    lib().orb_sort_pairs(k.ctypes.data, v.ctypes.data, k.size)
    return k, v
