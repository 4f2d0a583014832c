"""CPU tests of the oracle (Oracle-B) itself: pinned scalar pieces, camera known answers from
SURVEY 8(c), pipeline invariants and the committed golden fixtures."""
import math
import os

import numpy as np
import pytest

from oracle import oracle as O

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def test_half_constants_and_conversion():
    L = O.lib()
    assert L.orb_float_to_half(0.99) == 0x3BEC and L.orb_half_to_float(0x3BEC) == 0.990234375
    assert L.orb_float_to_half(0.01) == 0x211F
    assert abs(L.orb_half_to_float(0x211F) - 0.010002136) < 1e-9
    rng = np.random.default_rng(1)
    x = np.concatenate([rng.standard_normal(2000).astype(np.float32) * 10, np.float32([0, 1e-8, 65504, 1e-5, 6.1e-5])])
    got = np.array([L.orb_float_to_half(float(v)) for v in x], np.uint16)
    assert np.array_equal(got, x.astype(np.float16).view(np.uint16))     # RNE, subnormals kept


def test_pinned_exp_accuracy():
    L = O.lib()
    q = np.linspace(0, 9.3, 4001, dtype=np.float32)
    got = np.array([L.orb_exp_neg_half(float(v)) for v in q])
    ref = np.exp(-0.5 * q.astype(np.float64))
    assert np.max(np.abs(got - ref) / ref) < 5e-7
    assert L.orb_exp_neg_half(0.0) == 1.0


def test_camera_known_answers():
    # focalX = focalY = 869.1169 @720p, 1303.6753 @1080p, 2607.3506 @4K
    for (w, h), f in (((1280, 720), 869.1169), ((1920, 1080), 1303.6753), ((3840, 2160), 2607.3506)):
        P = O.perspective(45.0, w, h)
        assert abs(P[0] * w * 0.5 - f) < 2e-3 and abs(P[5] * h * 0.5 - f) < 2e-3
        assert P[11] == -1.0 and P[15] == 0.0
    # zNDC(zFront): -1 @0.1, ~0 @0.19998, 0.80018 @1, 0.960196 @5, 1 @1000
    P = O.perspective(45.0, 1280, 720)
    for z, want in ((0.1, -1.0), (0.19998, 0.0), (1.0, 0.80018), (5.0, 0.960196), (1000.0, 1.0)):
        ndc = (P[10] * (-z) + P[14]) / z
        assert abs(ndc - want) < 2e-4
    V = O.look_at((0, 2, 5), (0, 0, 0), (0, -1, 0))
    eye_view = V.reshape(4, 4).T @ np.array([0, 2, 5, 1], np.float32)
    assert np.allclose(eye_view[:3], 0, atol=1e-6)
    origin_view = V.reshape(4, 4).T @ np.array([0, 0, 0, 1], np.float32)
    assert abs(-origin_view[2] - math.sqrt(29.0)) < 1e-5          # zFront of the target


def test_product_camera_matches_oracle_bitwise():
    from gaussian_splat_b200 import camera
    for (w, h) in ((1280, 720), (1920, 1080), (3840, 2160), (333, 217)):
        for k in range(8):
            t = 2 * math.pi * k / 8
            eye = (5 * math.sin(t), 2.0, 5 * math.cos(t))
            V1, P1 = camera.app_camera(w, h, eye=eye)
            V2, P2 = O.app_camera(w, h, eye=eye)
            assert V1.tobytes() == V2.tobytes() and P1.tobytes() == P2.tobytes()


def test_pipeline_invariants(small_scene):
    w, h = 640, 360
    V, P = O.app_camera(w, h)
    for mode in (O.MODE_TILE, O.MODE_LAYER50):
        F = O.Frame(small_scene, V, P, w, h, mode=mode)
        t = F.tiles_touched.astype(np.int64)
        assert F.D == t.sum()
        assert np.array_equal(F.offsets, np.concatenate([[0], np.cumsum(t)[:-1]]).astype(np.uint32))
        ks, vs = F.keys_sorted, F.vals_sorted
        assert np.all(ks[1:] >= ks[:-1])
        order = np.argsort(F.keys_unsorted, kind="stable")
        assert np.array_equal(ks, F.keys_unsorted[order]) and np.array_equal(vs, F.vals_unsorted[order])
        tiles = (ks >> np.uint64(32)).astype(np.int64)
        assert tiles.max() < F.T
        r = F.tile_ranges
        cnt = np.bincount(tiles, minlength=F.T)
        assert np.array_equal((r[:, 1] - r[:, 0]).astype(np.int64), cnt)
        nz = cnt > 0
        assert np.array_equal(r[nz, 0].astype(np.int64), (np.cumsum(cnt) - cnt)[nz])
        rec = F.records
        vis = t > 0
        assert np.all(rec["z_front"][vis] >= 0.19) and np.all(rec["z_front"][~vis] == 0)
        if mode == O.MODE_TILE:
            assert np.array_equal((ks & np.uint64(0xFFFFFFFF)).astype(np.uint32), rec["bd"][vs] >> 16)
        else:
            assert np.array_equal((ks & np.uint64(0xFFFFFFFF)).astype(np.uint32), vs)


def test_reference_order_reverses_depth_key(small_scene):
    w, h = 320, 200
    V, P = O.app_camera(w, h)
    a = O.Frame(small_scene, V, P, w, h, order=O.ORDER_NEAR_FIRST)
    b = O.Frame(small_scene, V, P, w, h, order=O.ORDER_REFERENCE)
    ka, kb = a.keys_unsorted, b.keys_unsorted
    assert np.array_equal(ka >> np.uint64(32), kb >> np.uint64(32))
    assert np.all((ka & np.uint64(0xFFFF)) + (kb & np.uint64(0xFFFF)) == 0xFFFF)


def test_bands_compose_to_full_frame(small_scene):
    """Tile-row bands (the multi-GPU partition) reproduce the single-device image exactly."""
    w, h = 333, 217
    V, P = O.app_camera(w, h)
    full, _ = O.Frame(small_scene, V, P, w, h).blend()
    fb = np.zeros_like(full)
    ty = (h + 15) // 16
    cuts = [0, 3, 4, 9, ty]
    D = 0
    for a, b in zip(cuts[:-1], cuts[1:]):
        F = O.Frame(small_scene, V, P, w, h, tile_row_begin=a, tile_row_end=b)
        F.blend(fb=fb); D += F.D
    assert np.array_equal(fb, full)
    assert D == O.Frame(small_scene, V, P, w, h).D


def test_threads_do_not_change_results(small_scene):
    w, h = 320, 200
    V, P = O.app_camera(w, h)
    O.set_threads(1); a = O.Frame(small_scene, V, P, w, h); ia, ta = a.blend(1)
    O.set_threads(4); b = O.Frame(small_scene, V, P, w, h); ib, tb = b.blend(4)
    O.set_threads(1)
    assert np.array_equal(a.keys_sorted, b.keys_sorted) and np.array_equal(ia, ib) and ta == tb


@pytest.mark.parametrize("name,mode,K", [("tile", 0, 32), ("layer50", 1, 50), ("mlab", 2, 1 << 30)])
def test_reference_shader_golden_images(name, mode, K):
    """Committed images of the REFERENCE's own shader statements (Oracle-A, generated where /root/reference
    exists by tests/golden/make_golden.py). Oracle-B, in the shaders' literal depth order, must agree on
    every pixel whose list stayed under the cap: north_star tolerance 1/255, and mostly bit-identical."""
    g = np.load(os.path.join(GOLDEN, "golden_small.npz"))
    w, h = int(g["w"]), int(g["h"])
    F = O.Frame(g["ora_instances"], g["view"], g["proj"], w, h, mode=mode, order=O.ORDER_REFERENCE)
    img, _ = F.blend()
    ref = g["ora_" + name + "_image"]; under = g["ora_" + name + "_arrivals"] <= K
    assert under.mean() > 0.95
    d = np.abs(O.half_bits_to_float(img) - O.half_bits_to_float(ref)).max(axis=2)
    assert d[under].max() <= 1.0 / 255.0
    assert (img == ref).all(axis=2)[under].mean() > 0.99


@pytest.mark.parametrize("name", ["tile_near", "tile_ref", "layer50_near", "layer50_ref", "mlab"])
def test_golden_fixtures(name):
    """Committed vectors (tests/golden/make_golden.py): guards the oracle against silent drift."""
    g = np.load(os.path.join(GOLDEN, "golden_small.npz"))
    inst = g["instances"]; w, h = int(g["w"]), int(g["h"])
    mode = O.MODE_LAYER50 if name.startswith("layer50") else (O.MODE_MLAB if name == "mlab" else O.MODE_TILE)
    order = O.ORDER_REFERENCE if name.endswith("ref") else O.ORDER_NEAR_FIRST
    F = O.Frame(inst, g["view"], g["proj"], w, h, mode=mode, order=order)
    assert np.array_equal(F.keys_sorted, g[name + "_keys"])
    assert np.array_equal(F.vals_sorted, g[name + "_vals"])
    assert np.array_equal(F.tile_ranges, g[name + "_ranges"])
    img, _ = F.blend()
    assert np.array_equal(img, g[name + "_image"])


def test_sh_extension_against_float64_evaluation():
    """D11: max(0, loader colour + SH degrees 1..3 on normalize(pos - camera)), standard 3DGS basis."""
    from gaussian_splat_b200 import scene
    inst, sh = scene.make_instances(3000, 0x5EED0001, sh_rest=True)
    w, h = 320, 200
    eye = np.array([3.0, 2.0, -4.0])
    V, P = O.app_camera(w, h, eye=tuple(eye))
    C1 = 0.4886025119029199
    C2 = [1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396]
    C3 = [-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154, -0.4570457994644658,
          1.445305721320277, -0.5900435899266435]
    d = inst[:, 7:10].astype(np.float64) - eye
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    x, y, z = d.T
    xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
    basis = np.stack([-C1 * y, C1 * z, -C1 * x,
                      C2[0] * xy, C2[1] * yz, C2[2] * (2 * zz - xx - yy), C2[3] * xz, C2[4] * (xx - yy),
                      C3[0] * y * (3 * xx - yy), C3[1] * xy * z, C3[2] * y * (4 * zz - xx - yy),
                      C3[3] * z * (2 * zz - 3 * xx - 3 * yy), C3[4] * x * (4 * zz - xx - yy), C3[5] * z * (xx - yy),
                      C3[6] * x * (xx - 3 * yy)], axis=1)
    coef = sh.astype(np.float64).reshape(-1, 3, 15)
    for degree, nk in ((1, 3), (2, 8), (3, 15)):
        F = O.Frame(inst, V, P, w, h, sh_rest=sh, sh_degree=degree)
        want = np.maximum(0.0, inst[:, 10:13] + np.einsum("nk,nck->nc", basis[:, :nk], coef[:, :, :nk]))
        rec = F.records; vis = F.tiles_touched > 0
        got = np.stack([(rec["rg"] & 0xFFFF), (rec["rg"] >> 16), (rec["bd"] & 0xFFFF)], axis=1).astype(np.uint16)
        got = got.view(np.float16).astype(np.float64)
        assert vis.sum() > 500
        assert np.abs(got[vis] - want[vis]).max() <= 1e-3 * max(1.0, want[vis].max())      # half rounding
