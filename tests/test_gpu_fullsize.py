"""Full-size runs (BASELINE.json config 2: 1M gaussians, 1920x1080) checked through size-independent properties,
plus an Oracle-B spot check of a few tile rows, and the largest viewport the key layout admits."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _frame(ctx, V, P, w, h, **kw):
    import torch
    from gaussian_splat_b200 import capi
    fb = torch.full((h, w, 4), 7.0, dtype=torch.float16, device="cuda:0")
    opts = capi.make_opts(flags=capi.FLAG_KEEP_UNSORTED, **kw)
    for _ in range(2):
        ctx.render(torch.cuda.current_stream().cuda_stream, fb.data_ptr(), V, P, w, h, opts)
        torch.cuda.synchronize()
        st = ctx.status()
        if not st.overflow:
            break
    assert not st.overflow
    return fb.cpu().numpy().view(np.uint16), st


@pytest.fixture(scope="module")
def scene_1m():
    from gaussian_splat_b200 import scene
    inst, _ = scene.make_instances(1_000_000, scene.CONFIG_SEEDS[2])
    return inst


@pytest.mark.parametrize("mode", [0, 1, 2])
def test_full_size_pipeline_properties(gpu_ctx, scene_1m, mode):
    from gaussian_splat_b200 import camera, capi
    w, h = 1920, 1080
    V, P = camera.app_camera(w, h)
    gpu_ctx.upload(scene_1m)
    img, st = _frame(gpu_ctx, V, P, w, h, mode=mode)
    D = int(st.n_pairs)
    touched = gpu_ctx.fetch(capi.FETCH_TILES_TOUCHED).astype(np.uint64)
    assert D == int(touched.sum()) and D > 1_000_000
    assert np.array_equal(gpu_ctx.fetch(capi.FETCH_OFFSETS), (np.cumsum(touched) - touched).astype(np.uint32))
    ku, vu = gpu_ctx.fetch(capi.FETCH_KEYS_UNSORTED), gpu_ctx.fetch(capi.FETCH_VALS_UNSORTED)
    ks, vs = gpu_ctx.fetch(capi.FETCH_KEYS_SORTED), gpu_ctx.fetch(capi.FETCH_VALS_SORTED)
    assert ku.shape == ks.shape == (D,)
    # emission order: splat index ascending, each splat's tiles row-major
    assert np.all(np.diff(vu.astype(np.int64)) >= 0)
    # sortedness + stability: keys ascending, equal keys keep their emission (= splat index) order
    dk = np.diff(ks.astype(np.int64))                       # keys are below 2^48
    assert np.all(dk >= 0)
    assert np.all(np.diff(vs.astype(np.int64))[dk == 0] > 0)
    # the sorted list is a permutation of the emitted one: checksum of checksums over (key, value)
    mix = lambda k, v: (k * np.uint64(0x9E3779B97F4A7C15) ^ (v.astype(np.uint64) * np.uint64(0xC2B2AE3D27D4EB4F)))
    with np.errstate(over="ignore"):
        assert np.bitwise_xor.reduce(mix(ku, vu)) == np.bitwise_xor.reduce(mix(ks, vs))
        assert np.add.reduce(mix(ku, vu)) == np.add.reduce(mix(ks, vs))
    assert np.array_equal(np.sort(ku), ks)
    # ranges partition [0, D) by tile
    r = gpu_ctx.fetch(capi.FETCH_TILE_RANGES).astype(np.int64)
    tiles = (ks >> np.uint64(32)).astype(np.int64)
    nz = r[:, 1] > r[:, 0]
    assert int((r[:, 1] - r[:, 0]).sum()) == D
    assert np.array_equal(np.nonzero(nz)[0], np.unique(tiles))
    assert np.array_equal(tiles[r[nz, 0]], np.nonzero(nz)[0]) and np.array_equal(tiles[r[nz, 1] - 1], np.nonzero(nz)[0])
    # idempotence, and three tile-row bands compose to the same frame bit for bit
    img2, _ = _frame(gpu_ctx, V, P, w, h, mode=mode)
    assert np.array_equal(img, img2)
    comp = np.zeros_like(img)
    for a, b in ((0, 20), (20, 47), (47, 68)):
        part, _ = _frame(gpu_ctx, V, P, w, h, mode=mode, tile_row_begin=a, tile_row_end=b)
        comp[a * 16:min(b * 16, h)] = part[a * 16:min(b * 16, h)]
    assert np.array_equal(comp, img)
    assert img[..., 3].view(np.float16).astype(np.float32).mean() > 0.05          # something was drawn


def test_full_size_rows_against_oracle(gpu_ctx, scene_1m):
    """Oracle-B on two tile rows of the full-size frame (all 1M splats preprocessed on the CPU): bit-exact."""
    from gaussian_splat_b200 import camera
    from oracle import oracle as O
    w, h = 1920, 1080
    V, P = camera.app_camera(w, h)
    gpu_ctx.upload(scene_1m)
    img, _ = _frame(gpu_ctx, V, P, w, h, mode=0)
    O.set_threads(8)
    try:
        F = O.Frame(scene_1m, V, P, w, h, tile_row_begin=33, tile_row_end=35)
    finally:
        O.set_threads(1)
    ref = np.zeros((h, w, 4), np.uint16)
    F.blend(n_threads=2, fb=ref)
    assert np.array_equal(img[33 * 16:35 * 16], ref[33 * 16:35 * 16])


def test_largest_viewport(gpu_ctx):
    """4096 x 4096 = 65536 tiles: 16 tile bits, the most the (tile | depth) key admits; one more column is refused."""
    import torch
    from gaussian_splat_b200 import capi, scene
    from oracle import oracle as O
    from test_gpu_parity import _check_all
    inst, _ = scene.make_instances(3000, 0x5EED0007)
    inst[:, 4:7] *= 8.0
    w = h = 4096
    V, P = O.app_camera(w, h)
    F = _check_all(gpu_ctx, inst, V, P, w, h, 0, 0)
    assert F.T == 65536 and F.D > 0
    fb = torch.zeros((16, 4112, 4), dtype=torch.float16, device="cuda:0")
    with pytest.raises(capi.SplatError):
        gpu_ctx.render(torch.cuda.current_stream().cuda_stream, fb.data_ptr(), *O.app_camera(4112, 4096), 4112, 4096, capi.make_opts())
