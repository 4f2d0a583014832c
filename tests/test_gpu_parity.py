"""GPU parity: every stage of the CUDA path, through the C ABI, against Oracle-B.

Bar: integer/index work bit-exact (tiles touched, offsets, keys, sorted pairs, ranges);
records bit-exact (fp32 under pinned arithmetic); image max|diff| <= 1/255 and PSNR >= 50 dB
(bit-exact is expected and reported).
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _render(ctx, inst, V, P, w, h, **kw):
    import torch
    from gaussian_splat_b200 import capi
    fb = torch.full((h, w, 4), 7.0, dtype=torch.float16, device="cuda:0")
    opts = capi.make_opts(flags=capi.FLAG_KEEP_UNSORTED, **kw)
    ctx.render(torch.cuda.current_stream().cuda_stream, fb.data_ptr(), V, P, w, h, opts)
    torch.cuda.synchronize()
    st = ctx.status()
    if st.overflow:      # buffers were grown by status(); render again
        ctx.render(torch.cuda.current_stream().cuda_stream, fb.data_ptr(), V, P, w, h, opts)
        torch.cuda.synchronize()
        st = ctx.status()
    assert not st.overflow
    return fb.cpu().numpy().view(np.uint16), st


def _psnr(a, b):
    mse = np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2)
    return 200.0 if mse == 0 else 10 * np.log10(1.0 / mse)


def _check_all(ctx, inst, V, P, w, h, mode, order, K=0, rows=(0, 0), sh=None, sh_degree=0):
    from gaussian_splat_b200 import capi
    from oracle import oracle as O
    ctx.upload(inst)
    if sh is not None:
        ctx.upload_sh(sh, 3)
    got, st = _render(ctx, inst, V, P, w, h, mode=mode, order=order, max_fragments=K,
                      tile_row_begin=rows[0], tile_row_end=rows[1], sh_degree=sh_degree)
    F = O.Frame(inst, V, P, w, h, mode=mode, order=order, max_fragments=K or None,
                tile_row_begin=rows[0], tile_row_end=rows[1], sh_rest=sh, sh_degree=sh_degree)
    assert st.n_pairs == F.D
    assert np.array_equal(ctx.fetch(capi.FETCH_TILES_TOUCHED), F.tiles_touched)
    assert np.array_equal(ctx.fetch(capi.FETCH_OFFSETS), F.offsets)
    rec = ctx.fetch(capi.FETCH_RECORDS)
    assert rec.tobytes() == F.records.tobytes()
    assert np.array_equal(ctx.fetch(capi.FETCH_KEYS_UNSORTED), F.keys_unsorted)
    assert np.array_equal(ctx.fetch(capi.FETCH_VALS_UNSORTED), F.vals_unsorted)
    assert np.array_equal(ctx.fetch(capi.FETCH_KEYS_SORTED), F.keys_sorted)
    assert np.array_equal(ctx.fetch(capi.FETCH_VALS_SORTED), F.vals_sorted)
    assert np.array_equal(ctx.fetch(capi.FETCH_TILE_RANGES), F.tile_ranges)
    ref = np.full((h, w, 4), np.float16(7.0).view(np.uint16), np.uint16)
    F.blend(fb=ref)
    a, b = O.half_bits_to_float(got), O.half_bits_to_float(ref)
    assert np.abs(a - b).max() <= 1.0 / 255.0          # north_star tolerance
    assert _psnr(a, b) >= 50.0
    assert np.array_equal(got, ref), "image expected bit-exact under pinned arithmetic"
    return F


@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("order", [0, 1])
def test_stages_and_image_small(gpu_ctx, small_scene, mode, order):
    from oracle import oracle as O
    w, h = 640, 360
    V, P = O.app_camera(w, h)
    F = _check_all(gpu_ctx, small_scene, V, P, w, h, mode, order)
    assert F.D > 10000


@pytest.mark.parametrize("K", [1, 24, 32, 50, 64])
def test_fragment_caps(gpu_ctx, small_scene, K):
    from oracle import oracle as O
    w, h = 320, 200
    V, P = O.app_camera(w, h, eye=(0.0, 1.0, 9.0))
    for mode in (0, 1):
        _check_all(gpu_ctx, small_scene, V, P, w, h, mode, 0, K=K)


@pytest.mark.parametrize("degree", [1, 2, 3])
def test_sh_view_dependent_colour(gpu_ctx, degree):
    """Extension D11: SH degree 1..3 evaluated in K1; records (colour halves) and image bit-exact."""
    from gaussian_splat_b200 import scene
    from oracle import oracle as O
    inst, sh = scene.make_instances(20000, 0x5EED0001, sh_rest=True)
    w, h = 480, 270
    V, P = O.app_camera(w, h, eye=(3.0, 2.0, -4.0))
    F = _check_all(gpu_ctx, inst, V, P, w, h, 0, 0, sh=sh, sh_degree=degree)
    F0 = O.Frame(inst, V, P, w, h)
    assert (F.records["rg"] != F0.records["rg"]).mean() > 0.3       # the extension really changes colours
    # degree 0 with SH uploaded is the reference-parity path, untouched
    _check_all(gpu_ctx, inst, V, P, w, h, 0, 0, sh=sh, sh_degree=0)


def test_ragged_viewport_and_band(gpu_ctx, small_scene):
    """Width/height not multiples of 16, and a tile-row band (multi-GPU partition)."""
    from oracle import oracle as O
    w, h = 333, 217
    V, P = O.app_camera(w, h)
    _check_all(gpu_ctx, small_scene, V, P, w, h, 0, 0)
    _check_all(gpu_ctx, small_scene, V, P, w, h, 0, 0, rows=(3, 9))
    _check_all(gpu_ctx, small_scene, V, P, w, h, 1, 0, rows=(5, 14))


def test_band_shape_dense_tiles_and_stale_bin_words(gpu_ctx):
    """The band shape of K1 / K2 (a band of at most half the tile rows): K1 leaves bin words only for splats that
    produced pairs plus a visibility bitmap, K2 walks the bitmaps 1024 visible splats at a time. A scene packed
    around the view axis puts far more than 1024 visible splats into every 2048-splat tile (the chunk loop), and
    band / whole frame / other band in turn leave stale bin words behind that the bitmaps must mask. All three
    modes; stages bit-exact as everywhere."""
    from gaussian_splat_b200 import scene
    from oracle import oracle as O
    w, h = 320, 256                                  # 20 x 16 tiles
    V, P = O.app_camera(w, h)
    inst, _ = scene.make_instances(5000, 0x5EED00B1)
    inst[:, 7:10] *= 0.12                            # every splat near the target: all visible, all in the middle rows
    for rows in ((4, 12), (0, 0), (6, 10), (0, 8)):
        F = _check_all(gpu_ctx, inst, V, P, w, h, 0, 0, rows=rows)
        if rows == (4, 12):
            vis = (F.tiles_touched.reshape(-1)[:4096].reshape(2, 2048) > 0).sum(axis=1)
            assert vis.min() > 1024, vis              # the case the chunk loop exists for
    _check_all(gpu_ctx, inst, V, P, w, h, 1, 0, K=50, rows=(4, 12))
    _check_all(gpu_ctx, inst, V, P, w, h, 2, 0, rows=(4, 12))


def test_empty_and_tiny_scenes(gpu_ctx):
    import torch
    from gaussian_splat_b200 import capi, scene
    from oracle import oracle as O
    w, h = 128, 96
    V, P = O.app_camera(w, h)
    # empty scene: the whole framebuffer is overwritten with zeros
    gpu_ctx.upload(np.zeros((0, 14), np.float32))
    fb = torch.full((h, w, 4), 3.0, dtype=torch.float16, device="cuda:0")
    gpu_ctx.render(torch.cuda.current_stream().cuda_stream, fb.data_ptr(), V, P, w, h, None)
    torch.cuda.synchronize()
    assert float(fb.abs().max()) == 0.0
    # one splat, then three
    for n in (1, 3, 1025):
        inst, _ = scene.make_instances(n, 0x77)
        inst[:, 7:10] *= 0.2
        inst[:, 4:7] = 0.3
        _check_all(gpu_ctx, inst, V, P, w, h, 0, 0)
    # everything behind the camera
    inst, _ = scene.make_instances(100, 0x78)
    inst[:, 9] = 50.0
    F = _check_all(gpu_ctx, inst, V, P, w, h, 0, 0)
    assert F.D == 0


def test_huge_splats_overflow_regrow(gpu_ctx):
    """Splats covering the whole screen: D >> initial capacity exercises the overflow path."""
    from gaussian_splat_b200 import scene
    from oracle import oracle as O
    w, h = 1024, 768
    V, P = O.app_camera(w, h)
    inst, _ = scene.make_instances(600, 0x79)
    inst[:, 4:7] = 3.0
    F = _check_all(gpu_ctx, inst, V, P, w, h, 0, 0)
    assert F.D > (1 << 20)


def test_sort_standalone(gpu_ctx):
    import torch
    from oracle import oracle as O
    g = torch.Generator(device="cpu"); g.manual_seed(5)
    for n, bits in ((1, 64), (4097, 64), (300001, 64), (100000, 48), (50000, 12)):
        k = torch.randint(0, 2 ** 62, (n,), generator=g, dtype=torch.int64) * 4 + torch.randint(0, 4, (n,), generator=g)
        if bits < 64:
            k = k & ((1 << bits) - 1)
        v = torch.arange(n, dtype=torch.int32)
        kd, vd = k.cuda(), v.cuda()
        ka, va = torch.empty_like(kd), torch.empty_like(vd)
        alt = gpu_ctx.sort_pairs(torch.cuda.current_stream().cuda_stream, kd.data_ptr(), vd.data_ptr(),
                                 ka.data_ptr(), va.data_ptr(), n, 0, bits)
        torch.cuda.synchronize()
        rk, rv = (ka, va) if alt else (kd, vd)
        ek, ev = O.sort_pairs(k.numpy().view(np.uint64), v.numpy().view(np.uint32))
        assert np.array_equal(rk.cpu().numpy().view(np.uint64), ek)
        assert np.array_equal(rv.cpu().numpy().view(np.uint32), ev)


def test_sort_standalone_32bit_keys(gpu_ctx):
    """The key width frames sort (tile << 16 | half depth). Sizes straddle the pair-tile size; duplicates test stability."""
    import torch
    from oracle import oracle as O
    rng = np.random.default_rng(9)
    for n, bits, dup in ((1, 32, False), (6143, 32, False), (6145, 32, True), (300001, 32, False), (200000, 29, True), (50000, 9, False)):
        k = rng.integers(0, 2 ** 32, n, dtype=np.uint64)
        if dup:
            k = ((k & 0x3FF) << 16) | ((k >> 40) & 3)      # few distinct keys, long runs of equal ones
        k = (k & ((1 << bits) - 1)).astype(np.uint32)
        v = np.arange(n, dtype=np.uint32)
        kd, vd = torch.from_numpy(k.view(np.int32)).cuda(), torch.from_numpy(v.view(np.int32)).cuda()
        ka, va = torch.empty_like(kd), torch.empty_like(vd)
        alt = gpu_ctx.sort_pairs(torch.cuda.current_stream().cuda_stream, kd.data_ptr(), vd.data_ptr(),
                                 ka.data_ptr(), va.data_ptr(), n, 0, bits, key_bits=32)
        torch.cuda.synchronize()
        rk, rv = (ka, va) if alt else (kd, vd)
        ek, ev = O.sort_pairs(k.astype(np.uint64), v)
        assert np.array_equal(rk.cpu().numpy().view(np.uint32).astype(np.uint64), ek)
        assert np.array_equal(rv.cpu().numpy().view(np.uint32), ev)


def test_render_host_matches_device(gpu_ctx, small_scene):
    from gaussian_splat_b200 import capi
    from oracle import oracle as O
    w, h = 640, 360
    V, P = O.app_camera(w, h)
    gpu_ctx.upload(small_scene)
    dev, _ = _render(gpu_ctx, small_scene, V, P, w, h, mode=0)
    host = np.zeros((h, w, 4), np.uint16)
    gpu_ctx.render_host(host.ctypes.data, V, P, w, h, capi.make_opts())
    assert np.array_equal(host, dev)


def test_render_batch_host_pipeline_matches_single_frames(gpu_ctx, small_scene):
    """Camera batch (config 4 shape): pipelined copies must not disturb any frame."""
    import math
    import torch
    from gaussian_splat_b200 import capi
    from oracle import oracle as O
    w, h, n = 400, 240, 7
    gpu_ctx.upload(small_scene)
    cams = [O.app_camera(w, h, eye=(5 * math.sin(2 * math.pi * k / n), 2.0, 5 * math.cos(2 * math.pi * k / n))) for k in range(n)]
    views = np.stack([c[0] for c in cams]); projs = np.stack([c[1] for c in cams])
    out = torch.zeros((n, h, w, 4), dtype=torch.float16).pin_memory()
    gpu_ctx.render_batch_host([out[k].data_ptr() for k in range(n)], views, projs, w, h, capi.make_opts())
    for k in range(n):
        one = np.zeros((h, w, 4), np.uint16)
        gpu_ctx.render_host(one.ctypes.data, views[k], projs[k], w, h, capi.make_opts())
        assert np.array_equal(out[k].numpy().view(np.uint16), one), k
    F = O.Frame(small_scene, views[3], projs[3], w, h)
    ref, _ = F.blend()
    assert np.array_equal(out[3].numpy().view(np.uint16), ref)


@pytest.mark.parametrize("mode", [0, 1])
def test_render_batch_two_frames_in_flight(gpu_ctx, small_scene, mode):
    """splat_render_batch alternates two frame sets on two internal streams; every frame must equal the
    single-frame render, and a later upload (which drops the second set) must still work."""
    import math
    import torch
    from gaussian_splat_b200 import capi
    from oracle import oracle as O
    w, h, n = 320, 200, 5
    gpu_ctx.upload(small_scene)
    cams = [O.app_camera(w, h, eye=(5 * math.sin(2 * math.pi * k / n), 1.0 + 0.3 * k, 5 * math.cos(2 * math.pi * k / n))) for k in range(n)]
    views = np.stack([c[0] for c in cams]); projs = np.stack([c[1] for c in cams])
    fbs = torch.zeros((n, h, w, 4), dtype=torch.float16, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    for rep in range(2):
        fbs.zero_()
        gpu_ctx.render_batch(s, [fbs[k].data_ptr() for k in range(n)], views, projs, w, h, capi.make_opts(mode=mode))
        got = fbs.cpu().numpy().view(np.uint16)       # ordered after the batch on the caller's stream
        for k in range(n):
            one, _ = _render(gpu_ctx, small_scene, views[k], projs[k], w, h, mode=mode)
            assert np.array_equal(got[k], one), (rep, k)
    F = O.Frame(small_scene, views[n - 1], projs[n - 1], w, h, mode=mode)
    ref, _ = F.blend()
    assert np.array_equal(got[n - 1], ref)
    gpu_ctx.upload(small_scene[: len(small_scene) // 2])
    gpu_ctx.render_batch(s, [fbs[k].data_ptr() for k in range(2)], views[:2], projs[:2], w, h, capi.make_opts(mode=mode))
    half, _ = _render(gpu_ctx, small_scene[: len(small_scene) // 2], views[1], projs[1], w, h, mode=mode)
    assert np.array_equal(fbs[1].cpu().numpy().view(np.uint16), half)


def test_mlab_mode_stages_and_image(gpu_ctx, small_scene):
    """the 6-layer K-buffer mode (shaders/gaussian_splat.metal), every stage bit-exact."""
    from oracle import oracle as O
    w, h = 640, 360
    V, P = O.app_camera(w, h)
    F = _check_all(gpu_ctx, small_scene, V, P, w, h, mode=2, order=0)
    assert F.D > 0
    _check_all(gpu_ctx, small_scene, V, P, 333, 217, mode=2, order=0, rows=(3, 9))      # ragged viewport + band
    _check_all(gpu_ctx, small_scene[:50], V, P, 16, 16, mode=2, order=0)               # one tile: no sort pass at all


@pytest.mark.parametrize("name,mode,K", [("tile", 0, 32), ("layer50", 1, 50), ("mlab", 2, 1 << 30)])
def test_reference_shader_golden_images(gpu_ctx, name, mode, K):
    """The CUDA path against committed images of the REFERENCE's own shader statements (Oracle-A; see
    tests/golden/make_golden.py): 1/255 on every pixel whose list stayed under the cap."""
    import os
    from oracle import oracle as O
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "golden_small.npz"))
    w, h = int(g["w"]), int(g["h"])
    gpu_ctx.upload(g["ora_instances"])
    got, _ = _render(gpu_ctx, g["ora_instances"], g["view"], g["proj"], w, h, mode=mode, order=1)
    ref = g["ora_" + name + "_image"]; under = g["ora_" + name + "_arrivals"] <= K
    d = np.abs(O.half_bits_to_float(got) - O.half_bits_to_float(ref)).max(axis=2)
    assert d[under].max() <= 1.0 / 255.0
    assert (got == ref).all(axis=2)[under].mean() > 0.99


@pytest.mark.parametrize("name,mode,order", [("tile_near", 0, 0), ("tile_ref", 0, 1), ("layer50_near", 1, 0), ("layer50_ref", 1, 1),
                                             ("mlab", 2, 0)])
def test_committed_golden_fixture(gpu_ctx, name, mode, order):
    """tests/golden/golden_small.npz (made by tests/golden/make_golden.py): lists overflow K and early
    termination triggers in this scene. Sorted pairs, ranges and image bit-exact."""
    import os
    from gaussian_splat_b200 import capi
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "golden_small.npz"))
    w, h = int(g["w"]), int(g["h"])
    gpu_ctx.upload(g["instances"])
    got, st = _render(gpu_ctx, g["instances"], g["view"], g["proj"], w, h, mode=mode, order=order)
    assert np.array_equal(gpu_ctx.fetch(capi.FETCH_KEYS_SORTED), g[name + "_keys"])
    assert np.array_equal(gpu_ctx.fetch(capi.FETCH_VALS_SORTED), g[name + "_vals"])
    assert np.array_equal(gpu_ctx.fetch(capi.FETCH_TILE_RANGES), g[name + "_ranges"])
    assert np.array_equal(got, g[name + "_image"])


def test_pack_bgra8_matches_unorm_conversion(gpu_ctx):
    """Presentation step: RGBA16F -> BGRA8Unorm, byte-exact against
    rint(clamp(x,0,1)*255) on every finite half, the infinities, NaN and negative zero; odd pixel count."""
    import torch
    w, h = 257, 255                                               # odd pixel count: exercises the tail pixel
    bits = (np.arange(w * h * 4, dtype=np.uint32) * 2654435761 >> 16).astype(np.uint16)   # all 65536 patterns occur
    bits[:65536] = np.arange(65536, dtype=np.uint16)
    x = bits.view(np.float16).astype(np.float32)
    with np.errstate(invalid="ignore"):
        u = np.rint(np.clip(np.nan_to_num(x, nan=0.0, posinf=1.0, neginf=0.0), 0.0, 1.0) * np.float32(255.0)).astype(np.uint8)
    want = u.reshape(h, w, 4)[..., [2, 1, 0, 3]]
    fb = torch.from_numpy(bits.view(np.int16).reshape(h, w, 4)).cuda()
    out = torch.zeros((h, w, 4), dtype=torch.uint8, device="cuda")
    gpu_ctx.pack_bgra8(torch.cuda.current_stream().cuda_stream, fb.data_ptr(), out.data_ptr(), w, h)
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy(), want)


def test_host_delivery_as_bgra8(gpu_ctx, small_scene):
    """SPLAT_FLAG_HOST_BGRA8: host frames arrive in the reference drawable's format, equal to packing the half4 frame."""
    import torch
    from gaussian_splat_b200 import capi
    from oracle import oracle as O
    w, h, n = 400, 240, 3
    gpu_ctx.upload(small_scene)
    cams = [O.app_camera(w, h, eye=(1.0 + k, 2.0, 5.0)) for k in range(n)]
    views = np.stack([c[0] for c in cams]); projs = np.stack([c[1] for c in cams])
    half = torch.zeros((n, h, w, 4), dtype=torch.float16).pin_memory()
    gpu_ctx.render_batch_host([half[k].data_ptr() for k in range(n)], views, projs, w, h, capi.make_opts())
    out = torch.zeros((n, h, w, 4), dtype=torch.uint8).pin_memory()
    gpu_ctx.render_batch_host([out[k].data_ptr() for k in range(n)], views, projs, w, h, capi.make_opts(flags=capi.FLAG_HOST_BGRA8))
    x = half.numpy().astype(np.float32)
    want = np.rint(np.clip(x, 0.0, 1.0) * np.float32(255.0)).astype(np.uint8)[..., [2, 1, 0, 3]]
    assert np.array_equal(out.numpy(), want) and want.any()
    one = np.zeros((h, w, 4), np.uint8)
    gpu_ctx.render_host(one.ctypes.data, views[1], projs[1], w, h, capi.make_opts(flags=capi.FLAG_HOST_BGRA8))
    assert np.array_equal(one, want[1])


def test_two_contexts_in_one_process(gpu_ctx, small_scene):
    """SURVEY 8(b) threading row: one context per device, several contexts per process. A second context (on the second
    device when there is one, else on device 0 again) renders and sorts while the first stays usable; kernels that need
    more dynamic shared memory than the default set the attribute per device, not once per process."""
    import torch
    from gaussian_splat_b200 import capi
    from oracle import oracle as O
    dev = 1 if torch.cuda.device_count() > 1 else 0
    w, h = 320, 200
    V, P = O.app_camera(w, h)
    F = O.Frame(small_scene, V, P, w, h)
    ref, _ = F.blend()
    other = capi.Context(dev)
    try:
        other.upload(small_scene)
        with torch.cuda.device(dev):
            fb = torch.zeros((h, w, 4), dtype=torch.float16, device="cuda:%d" % dev)
            other.render(torch.cuda.current_stream().cuda_stream, fb.data_ptr(), V, P, w, h, capi.make_opts())
            torch.cuda.synchronize()
            assert np.array_equal(fb.cpu().numpy().view(np.uint16), ref)
            n = 100_000
            k = torch.randint(0, 2 ** 62, (n,), dtype=torch.int64, device="cuda:%d" % dev)
            v = torch.arange(n, dtype=torch.int32, device="cuda:%d" % dev)
            ka, va = torch.empty_like(k), torch.empty_like(v)
            alt = other.sort_pairs(torch.cuda.current_stream().cuda_stream, k.data_ptr(), v.data_ptr(), ka.data_ptr(), va.data_ptr(), n, 0, 64)
            torch.cuda.synchronize()
            rk = (ka if alt else k).cpu().numpy()
            assert np.all(np.diff(rk) >= 0)
        gpu_ctx.upload(small_scene)
        got, _ = _render(gpu_ctx, small_scene, V, P, w, h, mode=0)
        assert np.array_equal(got, ref)
    finally:
        other.close()


def test_status_reports_block_culling_and_chunks(gpu_ctx, small_scene):
    """splat_status: live blocks of the Morton-ordered scene (a tile-row band keeps fewer than the whole frame), parking
    slots >= pairs, the longest tile list."""
    from gaussian_splat_b200 import capi
    from oracle import oracle as O
    w, h = 640, 360
    V, P = O.app_camera(w, h)
    gpu_ctx.upload(small_scene)
    _, full = _render(gpu_ctx, small_scene, V, P, w, h, mode=0)
    _, band = _render(gpu_ctx, small_scene, V, P, w, h, mode=0, tile_row_begin=2, tile_row_end=5)
    n_blocks = (len(small_scene) + 63) // 64               # SCENE_BLOCK = 64 splats per culling block
    assert 0 < band.n_live_blocks < full.n_live_blocks <= n_blocks
    assert full.n_candidates >= full.n_pairs and full.n_candidates % 256 == 0
    r = gpu_ctx.fetch(capi.FETCH_TILE_RANGES).astype(np.int64)
    assert band.max_tile_pairs == int((r[:, 1] - r[:, 0]).max())


@pytest.mark.parametrize("case", ["orbit", "needles", "near"])
@pytest.mark.parametrize("name,mode,K", [("tile", 0, 32), ("layer50", 1, 50)])
def test_reference_shader_golden_images_vertex_stage_cases(gpu_ctx, case, name, mode, K):
    """More committed Oracle-A images (tests/golden/make_golden.py): an orbit view on a ragged viewport, 20:1 needles seen
    along an axis (the |b| <= 1e-8 eigenvector branch, gaussian_splat_tile.metal:76-80), a camera inside the cloud (the
    clip volume's near face). The CUDA path agrees with the reference's shader statements wherever the list stayed under
    the cap, and is bit-identical to Oracle-B on the whole frame."""
    import os
    from oracle import oracle as O
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "golden_small.npz"))
    inst, V, P = g[f"ora2_{case}_instances"], g[f"ora2_{case}_view"], g[f"ora2_{case}_proj"]
    w, h = (int(x) for x in g[f"ora2_{case}_wh"])
    gpu_ctx.upload(inst)
    got, _ = _render(gpu_ctx, inst, V, P, w, h, mode=mode, order=1)
    ref = g[f"ora2_{case}_{name}_image"]; under = g[f"ora2_{case}_{name}_arrivals"] <= K
    a, b = O.half_bits_to_float(got), O.half_bits_to_float(ref)
    tol = np.maximum(1.0 / 255.0, 8.0 * np.exp2(np.floor(np.log2(np.maximum(np.maximum(np.abs(a), np.abs(b)), 2.0 ** -14)))) * 2.0 ** -10)
    bad = (np.abs(a - b) > tol).any(axis=2) & under
    assert bad.sum() <= max(6, 5e-4 * w * h)                 # coverage ties only (enumerated on the CPU side of the test suite)
    assert (got == ref).all(axis=2)[under].mean() > 0.9
    F = O.Frame(inst, V, P, w, h, mode=mode, order=1)
    assert np.array_equal(got, F.blend()[0])


def test_render_into_a_peer_framebuffer(gpu_ctx, small_scene):
    """splat_peer_alloc: the framebuffer other ranks map and write their bands into (the fused exchange of the banded frame).
    On one GPU: a frame composed of two bands rendered straight into such a buffer equals the frame rendered into an ordinary
    device buffer, and the exported handle is 64 bytes. (Mapping it from another process needs a second GPU process:
    bench.py --workload c3 asserts that path against the one-GPU frame at N = 2, 4, 8.)"""
    import torch
    from gaussian_splat_b200 import capi, multigpu
    from oracle import oracle as O
    w, h = 640, 360
    V, P = O.app_camera(w, h)
    gpu_ctx.upload(small_scene)
    want, _ = _render(gpu_ctx, small_scene, V, P, w, h, mode=0)
    ptr, handle = capi.peer_alloc(w * h * 8)
    try:
        assert len(handle) == 64 and any(handle)
        s = torch.cuda.current_stream().cuda_stream
        for a, b in ((0, 9), (9, 23)):
            gpu_ctx.render(s, ptr, V, P, w, h, capi.make_opts(tile_row_begin=a, tile_row_end=b))
        torch.cuda.synchronize()
        got = torch.as_tensor(multigpu._DevView(ptr, (h, w, 4), "<f2"), device="cuda:0").cpu().numpy().view(np.uint16)
        assert np.array_equal(got, want)
    finally:
        capi.peer_free(ptr)


@pytest.mark.parametrize("mode", [0, 1, 2])
def test_fragment_near_discard_with_a_tiny_near_plane(gpu_ctx, mode):
    """The shaders' fragment stage discards depth < 0.001 (gaussian_splat_tile.metal:187, gaussian_splat_50layer.metal:154,
    gaussian_splat.metal:209). The app's near plane (zNDC = 0 at zFront = 0.19998) hides that rule; a caller's projection with
    a near plane of 1e-5 does not: splats with 1e-4 <= zFront < 1e-3 pass the vertex stage's cull and the clip volume, and
    must still draw nothing. Stages and image bit-exact against Oracle-B, which carries the same rule."""
    from oracle import oracle as O
    w, h = 256, 192
    V = O.look_at((0.0, 0.0, 0.0), (0.0, 0.0, -1.0), (0.0, -1.0, 0.0))
    P = O.perspective(45.0, w, h, 1e-5, 10.0)
    rng = np.random.default_rng(7)
    n = 400
    inst = np.zeros((n, 14), np.float32)
    inst[:, 0] = 1.0                                               # identity rotation
    z = np.concatenate([rng.uniform(2e-4, 9e-4, n // 2), rng.uniform(1.2e-3, 6e-3, n - n // 2)]).astype(np.float32)
    inst[:, 9] = -z                                                # straight ahead, zFront = z
    inst[:, 7:9] = (rng.uniform(-0.2, 0.2, (n, 2)) * z[:, None]).astype(np.float32)
    inst[:, 4:7] = (z[:, None] * rng.uniform(0.004, 0.02, (n, 3))).astype(np.float32)     # a few pixels on screen at any depth
    inst[:, 10:13] = rng.uniform(0.1, 1.0, (n, 3)).astype(np.float32)
    inst[:, 13] = rng.uniform(0.2, 0.9, n).astype(np.float32)
    F = _check_all(gpu_ctx, inst, V, P, w, h, mode, 0)
    touched = F.tiles_touched
    assert not touched[: n // 2].any()                             # zFront in [2e-4, 9e-4]: discarded although inside the clip volume
    assert (touched[n // 2:] > 0).sum() > n // 4                   # zFront > 1.2e-3: drawn


def _wall_scene(n, seed, spread=1.2, scale=0.25):
    """n splats on the plane z = 0 seen head-on from (0, 0, 5): every zFront is exactly 5, so every pair of a tile has the
    same half depth - the sorted order is decided by the instance index alone."""
    rng = np.random.default_rng(seed)
    inst = np.zeros((n, 14), np.float32)
    inst[:, 0] = 1.0
    inst[:, 4:7] = scale
    inst[:, 7:9] = rng.uniform(-spread, spread, (n, 2)).astype(np.float32)
    inst[:, 10:13] = rng.uniform(0.05, 1.0, (n, 3)).astype(np.float32)
    inst[:, 13] = rng.uniform(0.02, 0.2, n).astype(np.float32)
    return inst


@pytest.mark.parametrize("mode", [0, 1, 2])
def test_tile_sort_long_runs_of_equal_depth(gpu_ctx, mode):
    """The per-tile sort runs its network on 32-bit keys (half depth << 16 | slot) and then orders runs of equal half depth by
    instance index; a run longer than 48 sends the list through the 64-bit network instead. A wall of splats at one depth makes
    every list ONE run of several hundred pairs. Stages and image bit-exact against Oracle-B."""
    from oracle import oracle as O
    w, h = 160, 128
    V, P = O.app_camera(w, h, eye=(0.0, 0.0, 5.0))
    inst = _wall_scene(700, 3)
    F = _check_all(gpu_ctx, inst, V, P, w, h, mode, 0)
    r = F.tile_ranges.astype(np.int64)
    assert (r[:, 1] - r[:, 0]).max() > 300                       # long lists ...
    assert len(np.unique(F.records["bd"][F.tiles_touched > 0] >> 16)) == 1      # ... of one half depth
    _check_all(gpu_ctx, inst, V, P, w, h, mode, 1)               # the shader's literal (descending) order too


@pytest.mark.parametrize("mode", [0, 1, 2])
def test_tile_sort_lists_beyond_the_register_network(gpu_ctx, mode):
    """Tile lists of more than 2048 pairs take the network run in global / shared memory by the tile's CTA (exact, slow, rare:
    the benchmark scenes' longest list is 1586). 2600 large splats over a 96 x 64 viewport put about 2600 pairs into every
    tile; random depths in one mode pass, one depth for all (ties broken by instance index over the whole list) in another."""
    from gaussian_splat_b200 import scene
    from oracle import oracle as O
    w, h = 96, 64
    V, P = O.app_camera(w, h, eye=(0.0, 0.0, 5.0))
    inst, _ = scene.make_instances(2600, 0x5EED0C11)
    inst[:, 7:10] *= 0.15
    inst[:, 4:7] = 0.6
    inst[:, 13] *= 0.05
    F = _check_all(gpu_ctx, inst, V, P, w, h, mode, 0)
    r = F.tile_ranges.astype(np.int64)
    assert (r[:, 1] - r[:, 0]).max() > 2048
    wall = _wall_scene(2600, 5, spread=0.5, scale=0.8)
    F = _check_all(gpu_ctx, wall, V, P, w, h, mode, 0)
    assert (F.tile_ranges[:, 1].astype(np.int64) - F.tile_ranges[:, 0]).max() > 2048


@pytest.mark.parametrize("mode", [0, 1])
def test_non_finite_splats_among_ordinary_ones(gpu_ctx, small_scene, mode):
    """The scene is stored in Morton order and culled per 64-splat block: non-finite positions sort last and are ignored by the
    block bounds, a block holding a non-finite or huge scale is never culled by its bound. Splats with NaN / infinite position,
    scale or rotation scattered among ordinary ones: every stage and the image bit-exact against Oracle-B (which culls
    them through its comparisons), in the whole frame and in a band."""
    from oracle import oracle as O
    w, h = 480, 270
    V, P = O.app_camera(w, h)
    inst = small_scene[:6000].copy()
    rng = np.random.default_rng(13)
    bad = rng.choice(len(inst), 300, replace=False)
    inst[bad[:60], 7] = np.nan                      # position x
    inst[bad[60:100], 8] = np.inf
    inst[bad[100:140], 9] = -np.inf
    inst[bad[140:180], 4] = np.inf                  # scale
    inst[bad[180:220], 5] = np.nan
    inst[bad[220:250], 0:4] = 0.0                   # zero quaternion: normalisation divides by zero
    inst[bad[275:300], 6] = 1e30                    # a finite but absurd scale
    with np.errstate(all="ignore"):
        _check_all(gpu_ctx, inst, V, P, w, h, mode, 0)
        _check_all(gpu_ctx, inst, V, P, w, h, mode, 0, rows=(4, 9))
