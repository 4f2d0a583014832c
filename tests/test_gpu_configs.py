"""Parity at the sizes that are benchmarked (BASELINE.json configs 2-5): for every config the whole of K1-K4 - pairs per
splat, sorted keys, sorted values, tile ranges - is compared with Oracle-B bit for bit, plus four image tile-row bands each
(top row, the row with the most pairs, the middle row, the ragged bottom row), and for config 3 one band per rank count
(2, 4, 8) exactly as a rank of `bench.py --workload c3` renders it. K1's screening and the block culling are approximate,
camera-, band- and resolution-dependent: a wrong cull would pass every self-consistency check, not this one.
Matches shaders/gaussian_splat_tile.metal:94-157,178-267 and gaussian_splat_50layer.metal:149-223 through Oracle-B."""
import math
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

#        n          w     h     mode  seed cfg  cameras (orbit view indices of 64, or None = the app camera)
CONFIGS = {
    "c2": (1_000_000, 1920, 1080, 0, 2, [None]),
    "c3": (6_000_000, 3840, 2160, 0, 3, [None]),
    "c4": (3_000_000, 1920, 1080, 0, 4, [0, 16, 32, 48]),
    "c5": (2_000_000, 1920, 1080, 1, 5, [None]),
}


def _camera(k, w, h):
    from gaussian_splat_b200 import camera
    if k is None:
        return camera.app_camera(w, h)
    t = 2.0 * math.pi * k / 64
    return camera.app_camera(w, h, eye=(5.0 * math.sin(t), 2.0, 5.0 * math.cos(t)))


def _render(ctx, V, P, w, h, **kw):
    import torch
    from gaussian_splat_b200 import capi
    fb = torch.full((h, w, 4), 7.0, dtype=torch.float16, device="cuda:0")
    opts = capi.make_opts(**kw)
    for _ in range(3):
        ctx.render(torch.cuda.current_stream().cuda_stream, fb.data_ptr(), V, P, w, h, opts)
        torch.cuda.synchronize()
        st = ctx.status()
        if not st.overflow:
            break
    assert not st.overflow
    return fb.cpu().numpy().view(np.uint16), st


def _compare(ctx, inst, V, P, w, h, mode, rows=(0, 0), image_rows=None):
    from gaussian_splat_b200 import capi
    from oracle import oracle as O
    cores = os.cpu_count() or 1
    img, st = _render(ctx, V, P, w, h, mode=mode, tile_row_begin=rows[0], tile_row_end=rows[1])
    O.set_threads(cores)
    try:
        F = O.Frame(inst, V, P, w, h, mode=mode, tile_row_begin=rows[0], tile_row_end=rows[1])
    finally:
        O.set_threads(1)
    assert int(st.n_pairs) == F.D
    assert np.array_equal(ctx.fetch(capi.FETCH_TILES_TOUCHED), F.tiles_touched)
    assert np.array_equal(ctx.fetch(capi.FETCH_KEYS_SORTED), F.keys_sorted)
    assert np.array_equal(ctx.fetch(capi.FETCH_VALS_SORTED), F.vals_sorted)
    ranges = F.tile_ranges
    assert np.array_equal(ctx.fetch(capi.FETCH_TILE_RANGES), ranges)
    tiles_x, tiles_y = (w + 15) // 16, (h + 15) // 16
    a, b = (0, tiles_y) if rows[1] <= 0 else rows
    if image_rows is None:
        per_row = (ranges[:, 1].astype(np.int64) - ranges[:, 0]).reshape(tiles_y, tiles_x).sum(axis=1)
        image_rows = sorted({a, a + int(np.argmax(per_row[a:b])), (a + b) // 2, b - 1})
    ref = np.zeros((h, w, 4), np.uint16)
    for r in image_rows:
        F.blend(n_threads=cores, fb=ref, rows=(r, r + 1))
        y0, y1 = r * 16, min(r * 16 + 16, h)
        assert np.array_equal(img[y0:y1], ref[y0:y1]), ("image row", r)
        assert ref[y0:y1].any() or per_row[r] == 0
    return F


@pytest.mark.parametrize("name", ["c2", "c3", "c4", "c5"])
def test_config_sized_frame_matches_oracle(gpu_ctx, name):
    from gaussian_splat_b200 import scene
    n, w, h, mode, cfg, cams = CONFIGS[name]
    inst, _ = scene.make_instances(n, scene.CONFIG_SEEDS[cfg])
    gpu_ctx.upload(inst)
    for k in cams:
        V, P = _camera(k, w, h)
        F = _compare(gpu_ctx, inst, V, P, w, h, mode)
        assert F.D > n          # a real frame: more pairs than splats at every config
    if name == "c3":
        # one band of a 2-, 4- and 8-way split, as a rank of the band + gather workload renders it
        tiles_y = (h + 15) // 16
        V, P = _camera(None, w, h)
        for world, r in ((2, 1), (4, 2), (8, 5)):
            a, b = tiles_y * r // world, tiles_y * (r + 1) // world
            _compare(gpu_ctx, inst, V, P, w, h, mode, rows=(a, b))
    gpu_ctx.upload(np.zeros((0, 14), np.float32))          # release the scene's device memory for the next module
