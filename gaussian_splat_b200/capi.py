"""ctypes binding of libsplat_cuda.so (include/splat_cuda.h).

The library is the product; this file only marshals pointers.  There is no CPU
fallback: if the CUDA library is missing or fails, every call raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# SPLAT_CUDA_LIB: another build of the same library (tuning variants, `make -C csrc variant NAME=...`)
LIB_PATH = os.environ.get("SPLAT_CUDA_LIB") or os.path.join(_HERE, "libsplat_cuda.so")

MODE_TILE, MODE_LAYER50, MODE_MLAB = 0, 1, 2
ORDER_NEAR_FIRST, ORDER_REFERENCE = 0, 1
FLAG_TIMING, FLAG_KEEP_UNSORTED, FLAG_HOST_BGRA8, FLAG_COUNT_WORK = 1, 2, 4, 8
STATIC_SORT_MAX_PAIRS = 2097152      # splat_cuda.h: SPLAT_STATIC_SORT_MAX_PAIRS
(FETCH_RECORDS, FETCH_TILES_TOUCHED, FETCH_OFFSETS, FETCH_KEYS_UNSORTED, FETCH_VALS_UNSORTED,
 FETCH_KEYS_SORTED, FETCH_VALS_SORTED, FETCH_TILE_RANGES, FETCH_DIGIT_HISTOGRAMS, FETCH_INSTANCES,
 FETCH_SH_REST) = range(11)

RECORD_DTYPE = np.dtype([
    ("cx", "<f4"), ("cy", "<f4"), ("ux", "<f4"), ("uy", "<f4"), ("vx", "<f4"), ("vy", "<f4"),
    ("opacity", "<f4"), ("rg", "<u4"), ("bd", "<u4"), ("z_front", "<f4"),
    ("rect_lo", "<u4"), ("rect_hi", "<u4")])

EXPORTS = [
    "splat_create", "splat_destroy", "splat_upload", "splat_upload_sh", "splat_render",
    "splat_render_host", "splat_render_batch", "splat_render_batch_host", "splat_pack_bgra8", "splat_upload_ply_rows", "splat_frame_status", "splat_debug_fetch",
    "splat_stage_times", "splat_sort_pairs_u64_u32", "splat_sort_pairs_u32_u32", "splat_last_error_string", "splat_version",
    "splat_band_rows", "splat_gather_bands", "splat_gather_bands_batch", "splat_sort_pairs_u64_u32_static",
    "splat_peer_alloc", "splat_peer_free", "splat_peer_open", "splat_peer_close",
]


class SplatOpts(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "mode", "max_fragments", "order", "sh_degree", "tile_row_begin", "tile_row_end", "flags", "reserved")]


class SplatStatus(C.Structure):
    _fields_ = [("n_splats", C.c_uint32), ("n_visible_unused", C.c_uint32), ("n_pairs", C.c_uint64),
                ("pair_capacity", C.c_uint64), ("overflow", C.c_uint32), ("sort_passes", C.c_uint32),
                ("n_candidates", C.c_uint64), ("max_tile_pairs", C.c_uint32), ("n_live_blocks", C.c_uint32), ("blend_warp_candidates", C.c_uint64)]


class SplatError(RuntimeError):
    pass


_LIB = None


def lib():
    """Load libsplat_cuda.so (fails loudly if it has not been built)."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise SplatError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                             "(there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        vp = C.c_void_p
        L.splat_create.argtypes = [C.c_int, C.POINTER(vp)]
        L.splat_destroy.argtypes = [vp]
        L.splat_upload.argtypes = [vp, vp, C.c_uint32]
        L.splat_upload_sh.argtypes = [vp, vp, C.c_uint32, C.c_int]
        L.splat_render.argtypes = [vp, vp, vp, vp, vp, C.c_int, C.c_int, C.POINTER(SplatOpts)]
        L.splat_render_host.argtypes = [vp, vp, vp, vp, C.c_int, C.c_int, C.POINTER(SplatOpts)]
        L.splat_render_batch.argtypes = [vp, vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.POINTER(SplatOpts)]
        L.splat_render_batch_host.argtypes = [vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.POINTER(SplatOpts)]
        L.splat_upload_ply_rows.argtypes = [vp, vp, C.c_uint64, C.c_uint32, vp, C.c_float, C.c_int, C.POINTER(C.c_uint32)]
        L.splat_pack_bgra8.argtypes = [vp, vp, vp, vp, C.c_int, C.c_int]
        L.splat_frame_status.argtypes = [vp, C.POINTER(SplatStatus)]
        L.splat_debug_fetch.argtypes = [vp, C.c_int, vp, C.POINTER(C.c_size_t)]
        L.splat_stage_times.argtypes = [vp, C.POINTER(C.c_float)]
        L.splat_sort_pairs_u64_u32.argtypes = [vp, vp, vp, vp, vp, vp, C.c_uint64, C.c_int, C.c_int,
                                               C.POINTER(C.c_int)]
        L.splat_sort_pairs_u32_u32.argtypes = L.splat_sort_pairs_u64_u32.argtypes
        L.splat_sort_pairs_u64_u32_static.argtypes = [vp, vp, vp, vp, vp, C.c_uint64, C.c_int, C.c_int, C.POINTER(C.c_int)]
        L.splat_band_rows.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_int32)]
        L.splat_gather_bands.argtypes = [vp, vp, vp, C.POINTER(C.c_int32), C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
        L.splat_gather_bands_batch.argtypes = [vp, vp, vp, C.c_int, C.POINTER(C.c_int32), C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
        L.splat_peer_alloc.argtypes = [C.c_size_t, C.POINTER(vp), vp]
        L.splat_peer_free.argtypes = [vp]
        L.splat_peer_open.argtypes = [vp, C.POINTER(vp)]
        L.splat_peer_close.argtypes = [vp]
        L.splat_last_error_string.restype = C.c_char_p
        L.splat_version.restype = C.c_int
        _LIB = L
    return _LIB


def _check(rc, what):
    if rc != 0:
        raise SplatError(f"{what} failed ({rc}): {lib().splat_last_error_string().decode()}")


def make_opts(mode=MODE_TILE, max_fragments=0, order=ORDER_NEAR_FIRST, sh_degree=0,
              tile_row_begin=0, tile_row_end=0, flags=0):
    return SplatOpts(mode, max_fragments, order, sh_degree, tile_row_begin, tile_row_end, flags, 0)


_FETCH_DTYPE = {FETCH_RECORDS: RECORD_DTYPE, FETCH_TILES_TOUCHED: np.uint32, FETCH_OFFSETS: np.uint32,
                FETCH_KEYS_UNSORTED: np.uint64, FETCH_VALS_UNSORTED: np.uint32, FETCH_KEYS_SORTED: np.uint64,
                FETCH_VALS_SORTED: np.uint32, FETCH_TILE_RANGES: np.uint32,
                FETCH_DIGIT_HISTOGRAMS: np.uint32, FETCH_INSTANCES: np.float32, FETCH_SH_REST: np.float32}


class Context:
    """One splat_ctx = one CUDA device."""

    def __init__(self, device: int = 0):
        self._h = C.c_void_p()
        _check(lib().splat_create(device, C.byref(self._h)), "splat_create")
        self.device = device
        self.n = 0

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            lib().splat_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def upload(self, instances: np.ndarray):
        a = np.ascontiguousarray(instances, np.float32).reshape(-1, 14)
        _check(lib().splat_upload(self._h, a.ctypes.data, a.shape[0]), "splat_upload")
        self.n = a.shape[0]

    def upload_sh(self, sh_rest: np.ndarray, degree: int):
        a = np.ascontiguousarray(sh_rest, np.float32).reshape(-1, 45)
        _check(lib().splat_upload_sh(self._h, a.ctypes.data, a.shape[0], degree), "splat_upload_sh")

    def render(self, stream_ptr: int, fb_dev_ptr: int, view, proj, w: int, h: int, opts: SplatOpts | None = None):
        v = np.ascontiguousarray(view, np.float32); p = np.ascontiguousarray(proj, np.float32)
        _check(lib().splat_render(self._h, stream_ptr, fb_dev_ptr, v.ctypes.data, p.ctypes.data, w, h,
                                  C.byref(opts) if opts is not None else None), "splat_render")

    def render_host(self, fb_host_ptr: int, view, proj, w: int, h: int, opts: SplatOpts | None = None):
        v = np.ascontiguousarray(view, np.float32); p = np.ascontiguousarray(proj, np.float32)
        _check(lib().splat_render_host(self._h, fb_host_ptr, v.ctypes.data, p.ctypes.data, w, h,
                                       C.byref(opts) if opts is not None else None), "splat_render_host")

    def render_batch(self, stream_ptr: int, fb_dev_ptrs, views, projs, w: int, h: int, opts=None):
        n = len(fb_dev_ptrs)
        arr = (C.c_void_p * n)(*fb_dev_ptrs)
        v = np.ascontiguousarray(views, np.float32).reshape(n, 16)
        p = np.ascontiguousarray(projs, np.float32).reshape(n, 16)
        _check(lib().splat_render_batch(self._h, stream_ptr, arr, v.ctypes.data, p.ctypes.data, n, w, h,
                                        C.byref(opts) if opts is not None else None), "splat_render_batch")

    def upload_ply_rows(self, rows: np.ndarray, slot_dst, crop: float = 5.0, sh_degree: int = 0) -> int:
        """GPU-side PLY decode: rows float32 [n, stride] raw properties, slot_dst int32 [stride]. Returns splats kept."""
        rows = np.ascontiguousarray(rows, np.float32)
        slots = np.ascontiguousarray(slot_dst, np.int32)
        n, stride = (rows.shape if rows.ndim == 2 else (0, slots.size))
        kept = C.c_uint32(0)
        _check(lib().splat_upload_ply_rows(self._h, rows.ctypes.data if n else None, n, stride, slots.ctypes.data, crop,
                                           sh_degree, C.byref(kept)), "splat_upload_ply_rows")
        return int(kept.value)

    def pack_bgra8(self, stream_ptr: int, fb_dev_ptr: int, bgra_dev_ptr: int, w: int, h: int):
        """RGBA16F framebuffer -> BGRA8Unorm (the reference drawable's format), on the device."""
        _check(lib().splat_pack_bgra8(self._h, stream_ptr, fb_dev_ptr, bgra_dev_ptr, w, h), "splat_pack_bgra8")

    def render_batch_host(self, fb_host_ptrs, views, projs, w: int, h: int, opts=None):
        """n views end to end into host framebuffers (pinned for overlap), copies pipelined behind rendering."""
        n = len(fb_host_ptrs)
        arr = (C.c_void_p * n)(*fb_host_ptrs)
        v = np.ascontiguousarray(views, np.float32).reshape(n, 16)
        p = np.ascontiguousarray(projs, np.float32).reshape(n, 16)
        _check(lib().splat_render_batch_host(self._h, arr, v.ctypes.data, p.ctypes.data, n, w, h,
                                             C.byref(opts) if opts is not None else None), "splat_render_batch_host")

    def status(self) -> SplatStatus:
        s = SplatStatus()
        _check(lib().splat_frame_status(self._h, C.byref(s)), "splat_frame_status")
        return s

    def fetch(self, what: int) -> np.ndarray:
        nbytes = C.c_size_t(0)
        _check(lib().splat_debug_fetch(self._h, what, None, C.byref(nbytes)), "splat_debug_fetch")
        dt = np.dtype(_FETCH_DTYPE[what])
        out = np.zeros(nbytes.value // dt.itemsize, dt)
        if nbytes.value:
            _check(lib().splat_debug_fetch(self._h, what, out.ctypes.data, C.byref(nbytes)), "splat_debug_fetch")
        if what == FETCH_DIGIT_HISTOGRAMS:
            return out.reshape(8, 256)
        if what == FETCH_INSTANCES:
            return out.reshape(-1, 14)
        if what == FETCH_SH_REST:
            return out.reshape(-1, 45)
        return out.reshape(-1, 2) if what == FETCH_TILE_RANGES else out

    def stage_times(self):
        ms = (C.c_float * 8)()
        _check(lib().splat_stage_times(self._h, ms), "splat_stage_times")
        return dict(zip(("preprocess", "scan_duplicate", "sort", "ranges", "blend", "total"), list(ms)[:6]))

    def sort_pairs(self, stream_ptr, keys_ptr, vals_ptr, keys_alt_ptr, vals_alt_ptr, n, begin_bit=0, end_bit=64,
                   key_bits=64):
        """Stand-alone pair sort; key_bits = 64 or 32 (the width frames sort). Returns 1 if the result is in *_alt."""
        res = C.c_int(0)
        fn = lib().splat_sort_pairs_u64_u32 if key_bits == 64 else lib().splat_sort_pairs_u32_u32
        _check(fn(self._h, stream_ptr, keys_ptr, vals_ptr, keys_alt_ptr, vals_alt_ptr,
                  n, begin_bit, min(end_bit, key_bits), C.byref(res)), "splat_sort_pairs")
        return res.value


def band_rows(tiles_y: int, world: int, weights=None):
    """splat_band_rows: contiguous bands of tile rows for `world` ranks (even split, or minimising the heaviest band)."""
    out = np.zeros(2 * world, np.int32)
    wp = None
    if weights is not None:
        wt = np.ascontiguousarray(weights, np.float64)
        if wt.shape != (tiles_y,):
            raise ValueError("weights must be one non-negative number per tile row")
        wp = wt.ctypes.data_as(C.POINTER(C.c_double))
    _check(lib().splat_band_rows(tiles_y, world, wp, out.ctypes.data_as(C.POINTER(C.c_int32))), "splat_band_rows")
    return [(int(out[2 * r]), int(out[2 * r + 1])) for r in range(world)]


def gather_bands(nccl_comm_ptr: int, stream_ptr: int, fb_dev_ptr: int, pixel_rows, w: int, bytes_per_pixel: int, rank: int,
                 world: int, dst: int = 0):
    """splat_gather_bands: grouped ncclSend / ncclRecv of the framebuffer bands to rank dst, on the given stream."""
    rows = np.ascontiguousarray(np.asarray(pixel_rows, np.int32).reshape(-1))
    _check(lib().splat_gather_bands(C.c_void_p(nccl_comm_ptr), C.c_void_p(stream_ptr), C.c_void_p(fb_dev_ptr),
                                    rows.ctypes.data_as(C.POINTER(C.c_int32)), w, bytes_per_pixel, rank, world, dst),
           "splat_gather_bands")


def gather_bands_batch(nccl_comm_ptr: int, stream_ptr: int, fb_dev_ptrs, pixel_rows, w: int, bytes_per_pixel: int, rank: int,
                       world: int, dst: int = 0):
    """splat_gather_bands_batch: the gather of several framebuffers of one partition inside one NCCL group."""
    rows = np.ascontiguousarray(np.asarray(pixel_rows, np.int32).reshape(-1))
    arr = (C.c_void_p * len(fb_dev_ptrs))(*fb_dev_ptrs)
    _check(lib().splat_gather_bands_batch(C.c_void_p(nccl_comm_ptr), C.c_void_p(stream_ptr), arr, len(fb_dev_ptrs),
                                          rows.ctypes.data_as(C.POINTER(C.c_int32)), w, bytes_per_pixel, rank, world, dst),
           "splat_gather_bands_batch")


def peer_alloc(nbytes: int):
    """splat_peer_alloc: (device pointer, 64-byte handle) of a buffer other processes can map with peer_open."""
    p = C.c_void_p(); h = C.create_string_buffer(64)
    _check(lib().splat_peer_alloc(nbytes, C.byref(p), h), "splat_peer_alloc")
    return int(p.value), bytes(h.raw)


def peer_open(handle: bytes) -> int:
    p = C.c_void_p(); h = C.create_string_buffer(handle, 64)
    _check(lib().splat_peer_open(h, C.byref(p)), "splat_peer_open")
    return int(p.value)


def peer_close(ptr: int):
    _check(lib().splat_peer_close(C.c_void_p(ptr)), "splat_peer_close")


def peer_free(ptr: int):
    _check(lib().splat_peer_free(C.c_void_p(ptr)), "splat_peer_free")
