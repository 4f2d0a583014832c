"""N>1 host logic on CPU: band partition, gloo world_size-2 gather of framebuffer bands, view sharding,
and (with Oracle-B standing in for the device) that bands compose to the single-GPU frame bit-for-bit."""
import os
import socket

import numpy as np
import pytest

from gaussian_splat_b200 import multigpu as M


def test_band_rows_even_and_weighted():
    assert M.band_rows(135, 8) == [(135 * r // 8, 135 * (r + 1) // 8) for r in range(8)]
    sizes = [e - b for b, e in M.band_rows(135, 8)]
    assert sum(sizes) == 135 and max(sizes) - min(sizes) <= 1
    w = np.ones(68); w[30:38] = 50.0                       # a dense blob in the middle rows
    bands = M.band_rows(68, 4, w)
    assert bands[0][0] == 0 and bands[-1][1] == 68 and all(bands[i][1] == bands[i + 1][0] for i in range(3))
    loads = [w[b:e].sum() for b, e in bands]
    even = [w[b:e].sum() for b, e in M.band_rows(68, 4)]
    assert max(loads) < max(even) and max(loads) <= 150.0 + 1e-6
    assert all(e > b for b, e in bands)
    # more ranks than rows: trailing ranks get empty bands, nothing is lost
    b3 = M.band_rows(3, 8, np.ones(3))
    assert len(b3) == 8 and sum(e - b for b, e in b3) == 3
    assert M.bands_to_pixel_rows([(0, 34), (34, 68)], 1080) == [(0, 544), (544, 1080)]
    assert M.shard_views(64, 3, 8) == list(range(24, 32))
    assert sorted(sum((M.shard_views(10, r, 4) for r in range(4)), [])) == list(range(10))


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, h, w, bands, out_path):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rows = M.bands_to_pixel_rows(bands, h)
    fb = torch.full((h, w, 4), -1.0, dtype=torch.float16)
    a, b = rows[rank]
    fb[a:b] = float(rank + 1)                               # "render" this rank's band
    M.gather_bands(fb, rows, rank, world, dst=0)
    dist.barrier()
    if rank == 0:
        np.save(out_path, fb.numpy())
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_gather_bands_gloo(tmp_path, world):
    import torch.multiprocessing as mp
    h, w = 100, 48
    bands = M.band_rows((h + 15) // 16, world, np.arange(1, (h + 15) // 16 + 1))
    out = str(tmp_path / "fb.npy")
    mp.spawn(_worker, args=(world, _free_port(), h, w, bands, out), nprocs=world, join=True)
    fb = np.load(out)
    rows = M.bands_to_pixel_rows(bands, h)
    for r, (a, b) in enumerate(rows):
        assert np.all(fb[a:b] == r + 1)
    assert not np.any(fb == -1.0)


def test_weighted_bands_compose_to_full_frame(small_scene):
    from oracle import oracle as O
    w, h = 320, 200
    V, P = O.app_camera(w, h)
    F = O.Frame(small_scene, V, P, w, h)
    full, _ = F.blend()
    tiles_x, tiles_y = (w + 15) // 16, (h + 15) // 16
    r = F.tile_ranges
    row_pairs = (r[:, 1] - r[:, 0]).reshape(tiles_y, tiles_x).sum(axis=1)
    bands = M.band_rows(tiles_y, 3, row_pairs)
    fb = np.zeros((h, w, 4), np.uint16)
    total = 0
    for b, e in bands:
        Fb = O.Frame(small_scene, V, P, w, h, tile_row_begin=b, tile_row_end=e)
        Fb.blend(fb=fb); total += Fb.D
    assert np.array_equal(fb, full) and total == F.D


def test_measured_time_rebalancing_converges():
    """reweight_rows + band_rows on a synthetic cost model (cost per row != pairs per row): the heaviest band
    approaches the ideal share within a few iterations and never gets worse than the pair-balanced start."""
    from gaussian_splat_b200 import multigpu
    rng = np.random.default_rng(3)
    rows, world = 135, 8
    pairs = rng.uniform(0.5, 1.5, rows) * np.linspace(0.6, 1.4, rows)
    true_cost = pairs * np.linspace(1.6, 0.7, rows)              # upper rows cost more per pair
    w = pairs.copy()
    bands = multigpu.band_rows(rows, world, w)
    worst0 = max(true_cost[b:e].sum() for b, e in bands)
    worst = worst0
    for _ in range(5):
        times = [true_cost[b:e].sum() for b, e in bands]
        w = multigpu.reweight_rows(w, bands, times)
        bands = multigpu.band_rows(rows, world, w)
        assert bands[0][0] == 0 and bands[-1][1] == rows and all(bands[k][1] == bands[k + 1][0] for k in range(world - 1))
        worst = max(true_cost[b:e].sum() for b, e in bands)
    ideal = true_cost.sum() / world
    assert worst <= worst0 + 1e-9
    assert worst <= 1.08 * ideal, (worst, ideal, worst0)


def _shared_worker(rank, world, port, h, w, bands, name, out_path):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rows = M.bands_to_pixel_rows(bands, h)
    frame = M.SharedHostFrame(name, h, w, rank, world, register=False)      # no CUDA here: plain shared mapping
    fb = torch.full((h, w, 4), float(rank + 1), dtype=torch.float16)        # this rank's "device" framebuffer
    frame.deliver(fb, rows[rank])
    dist.barrier()
    if rank == world - 1:                                                   # any rank sees the whole frame
        np.save(out_path, frame.tensor.numpy().copy())
    dist.barrier()
    frame.close()
    dist.destroy_process_group()


def test_shared_host_frame_composes_bands_without_a_collective(tmp_path):
    """World-size-2 host delivery: every rank writes its band straight into one shared host frame."""
    import torch.multiprocessing as mp
    h, w, world = 100, 48, 2
    bands = M.band_rows((h + 15) // 16, world, np.arange(1, (h + 15) // 16 + 1))
    out = str(tmp_path / "frame.npy"); name = "splat_test_frame_%d" % os.getpid()
    mp.spawn(_shared_worker, args=(world, _free_port(), h, w, bands, name, out), nprocs=world, join=True)
    fr = np.load(out)
    for r, (a, b) in enumerate(M.bands_to_pixel_rows(bands, h)):
        assert np.all(fr[a:b] == r + 1)
    assert not os.path.exists(os.path.join("/dev/shm", name))


def test_band_rows_c_entry_matches_properties():
    """splat_band_rows (C ABI): contiguous, covering, heaviest band minimal against brute force on small cases."""
    import itertools
    import numpy as np
    from gaussian_splat_b200 import capi
    rng = np.random.default_rng(3)
    for tiles_y, world in ((7, 3), (9, 4), (12, 2), (5, 5), (3, 5)):
        wt = rng.integers(0, 50, tiles_y).astype(np.float64)
        bands = capi.band_rows(tiles_y, world, wt)
        assert len(bands) == world and bands[0][0] == 0 and max(b[1] for b in bands) == tiles_y
        flat = [x for b in bands if b[1] > b[0] for x in b]
        assert all(flat[i] <= flat[i + 1] for i in range(len(flat) - 1))
        assert sum(b[1] - b[0] for b in bands) == tiles_y
        heaviest = max(wt[a:b].sum() for a, b in bands)
        best = min(max(wt[a:b].sum() for a, b in zip((0,) + c, c + (tiles_y,)))
                   for k in range(0, min(world, tiles_y)) for c in itertools.combinations(range(1, tiles_y), k))
        assert heaviest <= best + 1e-6
    assert capi.band_rows(135, 8) == [(135 * r // 8, 135 * (r + 1) // 8) for r in range(8)]
