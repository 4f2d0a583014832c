"""Synthetic scene family: determinism, chunk independence, PLY layout."""
import numpy as np

from gaussian_splat_b200 import scene


def test_chunks_are_independent_of_chunking():
    a = scene.gen_chunk(10000, 7, 0, 3000)
    b = np.concatenate([scene.gen_chunk(10000, 7, 0, 1000), scene.gen_chunk(10000, 7, 1000, 2000)])
    assert a.tobytes() == b.tobytes()
    assert a.shape == (3000, 62)


def test_instances_layout_and_ranges():
    inst, sh = scene.make_instances(5000, 0x5EED0001, sh_rest=True)
    assert inst.shape == (5000, 14) and sh.shape == (5000, 45)
    assert np.all(np.abs(inst[:, 7:10]) < 4.9001)                 # D12: crop is a no-op
    assert np.all((inst[:, 13] > 0) & (inst[:, 13] < 1))          # sigmoid(opacity)
    assert np.all(inst[:, 4:7] > 0)                               # exp(scale)
    assert np.all((inst[:, 10:13] >= 0) & (inst[:, 10:13] <= 1))  # clamped DC colour
    again, _ = scene.make_instances(5000, 0x5EED0001)
    assert again.tobytes() == inst.tobytes()


def test_ply_file_layout(tmp_path):
    p = tmp_path / "s.ply"
    scene.write_ply(str(p), 100, 3)
    raw = open(p, "rb").read()
    head, body = raw.split(b"end_header\n", 1)
    lines = head.decode().splitlines()
    assert lines[0] == "ply" and lines[1] == "format binary_little_endian 1.0" and lines[2] == "element vertex 100"
    props = [l.split()[-1] for l in lines[3:]]
    assert props == scene.PROPS_FULL and len(props) == 62
    assert len(body) == 100 * 62 * 4
    assert np.frombuffer(body, "<f4").reshape(100, 62).tobytes() == scene.gen_chunk(100, 3, 0, 100).tobytes()
