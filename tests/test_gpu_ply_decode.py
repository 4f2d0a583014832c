"""GPU-side PLY decode against the CPU loader, which is itself pinned bit for bit to
the reference's src/ply_loader.cpp (tests/test_oracle_a_cpu.py). Bar: every float of every kept splat
bit-identical, same splats kept, same order."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

NAMES = (["x", "y", "z", "nx", "ny", "nz", "f_dc_0", "f_dc_1", "f_dc_2"] + [f"f_rest_{k}" for k in range(45)] +
         ["opacity", "scale_0", "scale_1", "scale_2", "rot_0", "rot_1", "rot_2", "rot_3"])
POINT_SLOT = {"x": 0, "y": 1, "z": 2, "nx": 3, "ny": 4, "nz": 5, "f_dc_0": 6, "f_dc_1": 7, "f_dc_2": 8, "opacity": 9,
              "scale_0": 10, "scale_1": 11, "scale_2": 12, "rot_0": 13, "rot_1": 14, "rot_2": 15, "rot_3": 16}
POINT_SLOT.update({f"f_rest_{k}": 17 + k for k in range(45)})


def _write(path, names, rows, ascii_=False):
    hdr = "ply\nformat %s 1.0\nelement vertex %d\n" % ("ascii" if ascii_ else "binary_little_endian", rows.shape[0])
    hdr += "".join(f"property float {p}\n" for p in names) + "end_header\n"
    with open(path, "wb") as f:
        f.write(hdr.encode("ascii"))
        if ascii_:
            for r in rows:
                f.write((" ".join(repr(float(v)) for v in r) + "\n").encode("ascii"))
        else:
            f.write(np.ascontiguousarray(rows, np.float32).tobytes())


def _expected(points62, crop=5.0):
    """The reference constructor's crop and field copy (src/instanced_splat_renderer.mm:360-388) on PointData rows."""
    p = points62
    with np.errstate(invalid="ignore"):
        keep = (np.abs(p[:, 0]) < crop) & (np.abs(p[:, 1]) < crop) & (np.abs(p[:, 2]) < crop)
    inst = np.concatenate([p[:, 13:17], p[:, 10:13], p[:, 0:3], p[:, 6:9], p[:, 9:10]], axis=1)
    return np.ascontiguousarray(inst[keep]), np.ascontiguousarray(p[keep, 17:62])


def _rows(n, seed):
    rng = np.random.default_rng(seed)
    rows = rng.normal(0, 1, (n, len(NAMES))).astype(np.float32)
    col = {p: i for i, p in enumerate(NAMES)}
    rows[:, :3] *= 3.0                                           # ~25% of the rows fall outside the +-5 crop
    rows[:, col["opacity"]] *= 6.0
    rows[:, col["scale_0"]:col["scale_2"] + 1] = rng.normal(-4, 3, (n, 3))
    # hard cases: exp overflow / underflow / subnormal results, infinities, NaN, all-zero DC, exact crop edge
    rows[0, col["scale_0"]] = 100.0; rows[1, col["scale_0"]] = -110.0; rows[2, col["scale_1"]] = -95.0
    rows[3, col["scale_2"]] = np.inf; rows[4, col["scale_2"]] = -np.inf; rows[5, col["scale_0"]] = np.nan
    rows[6, col["opacity"]] = 200.0; rows[7, col["opacity"]] = -200.0; rows[8, col["opacity"]] = np.nan
    rows[9, col["f_dc_0"]:col["f_dc_2"] + 1] = 0.0
    rows[10, 0] = 5.0; rows[11, 0] = np.nextafter(np.float32(5.0), np.float32(0)); rows[12, 1] = np.nan; rows[13, 2] = -np.inf
    rows[14, col["scale_1"]] = 88.72283935546875; rows[15, col["scale_1"]] = -103.97208404541016     # just past the thresholds
    rows[:16, :3] = np.where(np.isin(np.arange(16), (10, 11, 12, 13))[:, None], rows[:16, :3], 0.25)
    return rows


def _bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def _same(got, want):
    """Bit-identical, except that a NaN only has to be a NaN: x86 propagates the operand's sign and payload
    through arithmetic, the device returns the canonical NaN (garbage in either way)."""
    got = np.ascontiguousarray(got, np.float32); want = np.ascontiguousarray(want, np.float32)
    if got.shape != want.shape:
        return False
    ok = (_bits(got) == _bits(want)) | (np.isnan(got) & np.isnan(want))
    if not ok.all():
        idx = np.argwhere(~ok)[:8]
        print("mismatches at", idx.tolist(), [(got[tuple(i)], want[tuple(i)]) for i in idx])
    return bool(ok.all())


@pytest.mark.parametrize("n", [1, 127, 129, 40000])
def test_decode_matches_host_loader_bitwise(gpu_ctx, tmp_path, n):
    from gaussian_splat_b200 import capi, host_api as H
    rows = _rows(max(n, 16), 0xD0 + n)[:n] if n >= 16 else _rows(16, 0xD0)[11:11 + n]
    path = str(tmp_path / "a.ply")
    _write(path, NAMES, rows)
    pts = H.load_ply(path)
    assert pts.shape == (n, 62)
    want_inst, want_sh = _expected(pts)
    kept = gpu_ctx.upload_ply_rows(rows, [POINT_SLOT[p] for p in NAMES], crop=5.0, sh_degree=3)
    assert kept == want_inst.shape[0]
    got = gpu_ctx.fetch(capi.FETCH_INSTANCES)
    assert _same(got, want_inst)
    if kept:
        assert _same(gpu_ctx.fetch(capi.FETCH_SH_REST), want_sh)
    if n >= 40000:
        assert 0.5 * n < kept < 0.95 * n


def test_decode_header_variants(gpu_ctx, tmp_path):
    """Shuffled property order, unknown and duplicate properties, missing ones (PointData defaults), no crop."""
    from gaussian_splat_b200 import capi, host_api as H
    rng = np.random.default_rng(7)
    names = ["junk_a", "rot_3", "z", "opacity", "x", "scale_1", "f_dc_2", "x", "f_rest_44", "confidence", "y", "f_rest_0"]
    rows = rng.normal(0, 2, (3000, len(names))).astype(np.float32)
    path = str(tmp_path / "v.ply")
    _write(path, names, rows)
    pts = H.load_ply(path)
    slots = [POINT_SLOT.get(p, -1) for p in names]
    for crop in (5.0, 0.0):
        want_inst, want_sh = _expected(pts, crop if crop > 0 else np.inf)
        kept = gpu_ctx.upload_ply_rows(rows, slots, crop=crop, sh_degree=3)
        assert kept == want_inst.shape[0] and (crop > 0 or kept == 3000)
        assert _same(gpu_ctx.fetch(capi.FETCH_INSTANCES), want_inst)
        assert _same(gpu_ctx.fetch(capi.FETCH_SH_REST), want_sh)
    # everything cropped away -> empty scene, render still clears the framebuffer
    rows[:, 4] = 9.0; rows[:, 7] = 9.0
    assert gpu_ctx.upload_ply_rows(rows, slots, crop=5.0) == 0
    assert gpu_ctx.fetch(capi.FETCH_INSTANCES).shape == (0, 14)


@pytest.mark.parametrize("ascii_", [False, True])
def test_renderer_with_device_decode_matches_host_decode(tmp_path, ascii_):
    """InstancedSplatRenderer(path, SplatDecode::Device): same instances, same image as the host-decoded renderer."""
    import torch
    from gaussian_splat_b200 import capi, host_api as H
    n, w, h = 6000, 320, 200
    rows = _rows(n, 0xA5)
    rows[:, NAMES.index("scale_0"):NAMES.index("scale_2") + 1] = np.random.default_rng(3).normal(-3.0, 0.5, (n, 3))
    rows[~np.isfinite(rows)] = 0.5
    path = str(tmp_path / ("t.ply"))
    _write(path, NAMES, rows, ascii_=ascii_)
    imgs = []
    cam = H.TrackballCamera(); cam.set_viewport(w, h); cam.set_position(0, 2, 9); cam.set_target(0, 0, 0)
    for dev in (False, True):
        r = H.InstancedSplatRenderer(path, device_decode=dev)
        assert r.initialize(0)
        fb = torch.zeros((h, w, 4), dtype=torch.float16, device="cuda:0")
        r.set_options(capi.make_opts(sh_degree=2))
        r.render(torch.cuda.current_stream().cuda_stream, fb.data_ptr(), cam.view(), cam.proj(), w, h)
        torch.cuda.synchronize()
        imgs.append((r.point_count(), r.instances().copy(), fb.cpu().numpy().view(np.uint16)))
    assert imgs[0][0] == imgs[1][0] and 0.5 * n < imgs[0][0] < n
    assert _same(imgs[1][1], imgs[0][1])
    assert np.array_equal(imgs[0][2], imgs[1][2]) and imgs[0][2].any()
