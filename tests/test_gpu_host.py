"""GPU: the C++ InstancedSplatRenderer (reference signatures) and the headless driver give the same
image as the C ABI called directly, and both match Oracle-B."""
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_cpp_renderer_matches_c_abi_and_oracle(tmp_path):
    import torch
    from gaussian_splat_b200 import capi, host_api as H, scene
    from oracle import oracle as O
    n, w, h = 20000, 640, 360
    p = str(tmp_path / "s.ply")
    scene.write_ply(p, n, 0x5EED0001)
    r = H.InstancedSplatRenderer(p)
    assert r.point_count() == n and r.initialize(0)
    cam = H.TrackballCamera(); cam.set_viewport(w, h); cam.set_position(0, 2, 5); cam.set_target(0, 0, 0)
    V, P = cam.view(), cam.proj()
    fb = torch.full((h, w, 4), 5.0, dtype=torch.float16, device="cuda:0")
    for mode in (capi.MODE_TILE, capi.MODE_LAYER50):
        r.set_options(capi.make_opts(mode=mode))
        r.render(torch.cuda.current_stream().cuda_stream, fb.data_ptr(), V, P, w, h)
        torch.cuda.synchronize()
        got = fb.cpu().numpy().view(np.uint16)
        F = O.Frame(r.instances(), V, P, w, h, mode=mode)
        ref, _ = F.blend()
        assert np.array_equal(got, ref)


def test_headless_driver_runs_and_writes_a_frame(tmp_path):
    from gaussian_splat_b200 import scene
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "gaussian_splat_b200", "splat_headless")
    p = str(tmp_path / "s.ply"); out = str(tmp_path / "f.ppm")
    scene.write_ply(p, 30000, 0x5EED0001)
    res = subprocess.run([exe, p, "--width", "320", "--height", "200", "--frames", "4", "--out", out],
                         capture_output=True, text=True, timeout=120)
    assert res.returncode == 0, res.stderr
    assert "points: 30000" in res.stdout and "frames/s" in res.stdout
    raw = open(out, "rb").read()
    assert raw.startswith(b"P6\n320 200\n255\n") and len(raw) == 15 + 320 * 200 * 3
    assert np.frombuffer(raw[15:], np.uint8).max() > 100           # something was drawn


def test_headless_driver_replays_an_interaction_trace(tmp_path):
    """mouse / scroll events through the reference's camera handlers, a frame per `frame` line;
    the final camera pose equals the one the same events give the host camera class directly."""
    from gaussian_splat_b200 import host_api as H, scene
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "gaussian_splat_b200", "splat_headless")
    p = str(tmp_path / "s.ply"); out = str(tmp_path / "t.ppm"); tr = str(tmp_path / "trace.txt")
    scene.write_ply(p, 20000, 0x5EED0001)
    events = [("frame",), ("down", 100.0, 80.0, 0), ("move", 140.0, 95.0), ("frame",), ("move", 180.0, 60.0), ("up",),
              ("scroll", 1.5), ("frame",), ("down", 50.0, 50.0, 1), ("move", 70.0, 90.0), ("up",), ("scroll", -0.5), ("frame",)]
    with open(tr, "w") as f:
        for e in events:
            f.write(" ".join(str(x) for x in e) + "\n")
    res = subprocess.run([exe, p, "--width", "320", "--height", "200", "--trace", tr, "--out", out],
                         capture_output=True, text=True, timeout=120)
    assert res.returncode == 0, res.stderr
    cam = H.TrackballCamera(); cam.set_viewport(320, 200); cam.set_target(0, 0, 0); cam.set_position(0, 2, 5)
    for e in events:
        if e[0] == "down": cam.mouse_down(e[1], e[2], e[3])
        elif e[0] == "move": cam.mouse_move(e[1], e[2])
        elif e[0] == "up": cam.mouse_up()
        elif e[0] == "scroll": cam.scroll(e[1])
    line = [ln for ln in res.stdout.splitlines() if ln.startswith("trace:")][0]
    assert line.startswith("trace: 4 frames")
    got = np.array([float(x) for x in line.split("camera at")[1].split()], np.float32)
    assert np.array_equal(got, np.asarray(cam.position(), np.float32))
    assert open(out, "rb").read().startswith(b"P6\n320 200\n255\n")


def test_cpp_renderer_defaults_are_what_the_reference_runs(tmp_path):
    """An unmodified caller of InstancedSplatRenderer gets the reference app's image: the 50-layer mode
    (src/instanced_splat_renderer.mm:38) with the shader's literal, descending depth order (gaussian_splat_50layer.metal:201)."""
    import torch
    from gaussian_splat_b200 import host_api as H, scene
    from oracle import oracle as O
    n, w, h = 20000, 480, 270
    p = str(tmp_path / "s.ply")
    scene.write_ply(p, n, 0x5EED0001)
    r = H.InstancedSplatRenderer(p)
    assert r.initialize(0)
    cam = H.TrackballCamera(); cam.set_viewport(w, h); cam.set_position(0, 2, 5); cam.set_target(0, 0, 0)
    fb = torch.full((h, w, 4), 5.0, dtype=torch.float16, device="cuda:0")
    r.render(torch.cuda.current_stream().cuda_stream, fb.data_ptr(), cam.view(), cam.proj(), w, h)      # no setOptions call
    torch.cuda.synchronize()
    F = O.Frame(r.instances(), cam.view(), cam.proj(), w, h, mode=O.MODE_LAYER50, order=O.ORDER_REFERENCE)
    ref, _ = F.blend()
    assert np.array_equal(fb.cpu().numpy().view(np.uint16), ref)


def test_cpp_renderer_recovers_from_pair_buffer_overflow(tmp_path):
    """A zoomed-in camera on large splats needs far more pairs than the initial capacity. The frame that overflows is
    rendered empty (never a truncated list); the next render() call learns of it without a synchronising copy, the
    library grows its buffers, the class enqueues the frame again, and that frame is complete - the reference never
    drops splats (src/instanced_splat_renderer.mm:496-499)."""
    import torch
    from gaussian_splat_b200 import capi, host_api as H, scene
    from oracle import oracle as O
    n, w, h = 600, 1024, 768
    raw = scene.gen_chunk(n, 0x79, 0, n, False).copy()
    for a in range(3):
        raw[:, scene.PROPS_NOSH.index("scale_%d" % a)] = np.float32(np.log(3.0))          # stored as log: scale 3.0, screen-filling splats
    p = str(tmp_path / "big.ply")
    with open(p, "wb") as f:
        f.write(("ply\nformat binary_little_endian 1.0\nelement vertex %d\n" % n +
                 "".join("property float %s\n" % q for q in scene.PROPS_NOSH) + "end_header\n").encode("ascii"))
        f.write(raw.astype("<f4").tobytes())
    r = H.InstancedSplatRenderer(p)
    assert r.initialize(0)
    r.set_options(capi.make_opts(mode=capi.MODE_TILE))
    cam = H.TrackballCamera(); cam.set_viewport(w, h); cam.set_position(0, 2, 5); cam.set_target(0, 0, 0)
    V, P = cam.view(), cam.proj()
    F = O.Frame(r.instances(), V, P, w, h)
    assert F.D > (1 << 20)                                   # more than the initial capacity
    ref, _ = F.blend(n_threads=os.cpu_count() or 1)
    fb = torch.full((h, w, 4), 5.0, dtype=torch.float16, device="cuda:0")
    s = torch.cuda.current_stream().cuda_stream
    r.render(s, fb.data_ptr(), V, P, w, h)
    torch.cuda.synchronize()
    assert not fb.cpu().numpy().any()                        # overflowed: empty, not truncated
    r.render(s, fb.data_ptr(), V, P, w, h)                   # sees the overflow, grows, renders again
    torch.cuda.synchronize()
    assert np.array_equal(fb.cpu().numpy().view(np.uint16), ref)
