"""Host layer that keeps the reference's C++ surface (gaussian_splat_b200/host): loader and camera
against the reference's own code compiled into oracle/_ref, renderer constructor semantics. No GPU."""
import math

import numpy as np
import pytest

from gaussian_splat_b200 import host_api as H
from gaussian_splat_b200 import scene
from oracle import oracle_a as A

needs_ref = pytest.mark.skipif(not (A.available("camera") and A.available("ply")), reason="oracle/_ref not built")


def _write(path, props, rows, fmt="binary_little_endian"):
    with open(path, "wb") as f:
        f.write((f"ply\nformat {fmt} 1.0\nelement vertex {len(rows)}\n" +
                 "".join(f"property float {q}\n" for q in props) + "end_header\n").encode())
        if fmt == "ascii":
            for r in rows:
                f.write((" ".join(repr(float(v)) for v in r) + "\n").encode())
        else:
            f.write(np.asarray(rows, "<f4").tobytes())


@needs_ref
def test_loader_matches_reference_loader_bitwise(tmp_path):
    p = str(tmp_path / "full.ply")
    scene.write_ply(p, 70000, 0x5EED0001)                      # > one decode thread's share
    a, b = H.load_ply(p), A.load_ply(p)
    assert a.shape == b.shape == (70000, 62) and a.tobytes() == b.tobytes()
    # subset of properties, shuffled order, an unknown property, all-zero DC quirk, out-of-range f_rest index
    props = ["scale_2", "x", "junk", "opacity", "f_dc_1", "z", "f_dc_0", "rot_3", "y", "f_dc_2", "f_rest_44", "f_rest_45", "scale_0"]
    rng = np.random.default_rng(3)
    rows = rng.standard_normal((257, len(props))).astype("<f4")
    rows[5, [4, 6, 9]] = 0.0
    p2 = str(tmp_path / "subset.ply")
    _write(p2, props, rows)
    a, b = H.load_ply(p2), A.load_ply(p2)
    assert a.tobytes() == b.tobytes()
    assert np.array_equal(a[5, 6:9], [0, 0, 0])
    # declared big-endian is read as little-endian by both (src/ply_loader.cpp:228)
    p3 = str(tmp_path / "be.ply")
    _write(p3, props, rows, fmt="binary_big_endian")
    assert H.load_ply(p3).tobytes() == A.load_ply(p3).tobytes()


def test_loader_errors_and_ascii(tmp_path):
    assert H.load_ply(str(tmp_path / "missing.ply")) is None
    (tmp_path / "bad.ply").write_bytes(b"plx\nend_header\n")
    assert H.load_ply(str(tmp_path / "bad.ply")) is None
    (tmp_path / "empty.ply").write_bytes(b"ply\nformat ascii 1.0\nelement vertex 0\nproperty float x\nend_header\n")
    assert H.load_ply(str(tmp_path / "empty.ply")) is None
    props = ["x", "y", "z", "f_dc_0", "f_dc_1", "f_dc_2", "opacity", "scale_0"]
    rows = [[1, 2, 3, 1, -1, 0, 0, -4], [4, 5, 6, 0, 0, 0, 2, 0]]
    _write(str(tmp_path / "a.ply"), props, rows, fmt="ascii")
    pts = H.load_ply(str(tmp_path / "a.ply"))
    assert pts.shape == (2, 62)                                  # the reference returns 4 here (resize + push_back)
    assert np.allclose(pts[0, 6:9], [0.78209, 0.21791, 0.5], atol=1e-5) and abs(pts[0, 9] - 0.5) < 1e-7
    assert abs(pts[0, 10] - math.exp(-4)) < 1e-8 and pts[0, 11] == np.float32(0.01)   # default scale kept


@needs_ref
def test_camera_matches_reference_camera_bitwise():
    def drive(c):
        out = []
        c.set_viewport(1280, 720); c.set_position(0, 2, 5); c.set_target(0, 0, 0)
        out += [c.view(), c.proj()]
        c.mouse_down(100, 120, 0); c.mouse_move(160, 90); c.mouse_move(161, 300); c.mouse_up()
        out += [c.view(), c.position()]
        c.mouse_down(10, 10, 2); c.mouse_move(-40, 77); c.mouse_up()
        out += [c.view(), c.position(), c.target()]
        c.scroll(3.0); out += [c.view()]
        c.scroll(-50.0); out += [c.position()]                   # clamps at maxDistance
        c.set_distance(0.01); out += [c.position()]              # clamps at minDistance
        c.mouse_move(5, 5); out += [c.view()]                    # no button held: only lastMousePos moves
        c.set_viewport(333, 217); out += [c.proj()]
        return out
    for a, b in zip(drive(H.TrackballCamera()), drive(A.RefCamera())):
        assert a.tobytes() == b.tobytes()


def test_renderer_constructor_crops_and_copies(tmp_path):
    n = 5000
    raw = scene.gen_chunk(n, 11, 0, n)
    raw[::7, 0] *= 3.0                                           # push some outside |x| < 5
    raw[3, 1] = 5.0                                              # exactly on the bound: dropped (strict <)
    p = str(tmp_path / "crop.ply")
    with open(p, "wb") as f:
        f.write(("ply\nformat binary_little_endian 1.0\nelement vertex %d\n" % n +
                 "".join(f"property float {q}\n" for q in scene.PROPS_FULL) + "end_header\n").encode())
        f.write(raw.tobytes())
    r = H.InstancedSplatRenderer(p)
    inst, _ = scene.raw_to_instances(raw)
    assert 0 < r.point_count() == inst.shape[0] < n
    got = r.instances()
    assert np.array_equal(got[:, :4], inst[:, :4]) and np.array_equal(got[:, 7:10], inst[:, 7:10])
    assert np.allclose(got, inst, rtol=3e-7, atol=1e-7)
    # a failed load leaves an empty renderer whose render() is a silent no-op (src/instanced_splat_renderer.mm:346-349,434-440)
    e = H.InstancedSplatRenderer(str(tmp_path / "nope.ply"))
    assert e.point_count() == 0
    e.render(None, None, np.eye(4, dtype=np.float32), np.eye(4, dtype=np.float32), 64, 64)
