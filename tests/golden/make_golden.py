"""Generates tests/golden/golden_small.npz from Oracle-B (run from the repo root:
`python tests/golden/make_golden.py`). The reference ships no golden vectors, so
these pin OUR restatement against drift; Oracle-A (oracle/_ref, the .metal text compiled as
C++) is the independent check of the restatement itself (tests/test_oracle_a.py)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from gaussian_splat_b200 import scene  # noqa: E402
from oracle import oracle as O  # noqa: E402

n, w, h = 1500, 160, 96
inst, _ = scene.make_instances(n, 0x60D)
inst[:, 4:7] *= 6.0          # larger splats so lists overflow K and early termination triggers
V, P = O.app_camera(w, h)
out = {"instances": inst, "view": V, "proj": P, "w": w, "h": h}
for name, mode, order in (("tile_near", 0, 0), ("tile_ref", 0, 1), ("layer50_near", 1, 0), ("layer50_ref", 1, 1),
                          ("mlab", 2, 0)):
    F = O.Frame(inst, V, P, w, h, mode=mode, order=order)
    img, _ = F.blend()
    out[name + "_keys"] = F.keys_sorted
    out[name + "_vals"] = F.vals_sorted
    out[name + "_ranges"] = F.tile_ranges
    out[name + "_image"] = img
    print(name, "D", F.D, "alpha mean", O.half_bits_to_float(img)[..., 3].mean())
# Images of the REFERENCE's own shader statements (Oracle-A: oracle/_ref/libref_metal.so, the .metal text
# compiled as C++ where it lies under /root/reference). They cannot be regenerated on the GPU box, so they
# are committed here; tests compare Oracle-B and the CUDA path with them under the documented tolerance.
from oracle import oracle_a as A  # noqa: E402
if A.available("metal"):
    # a sparser variant of the scene: most per-pixel lists stay under the caps, beyond which TILE differs
    # from the shader text by design
    inst_a, _ = scene.make_instances(n, 0x60D)
    inst_a[:, 4:7] *= 2.5
    out["ora_instances"] = inst_a
    for name, mode in (("tile", 0), ("layer50", 1), ("mlab", 2)):
        img, arrivals = A.render(inst_a, V, P, w, h, mode)
        out["ora_" + name + "_image"] = img
        out["ora_" + name + "_arrivals"] = arrivals
        print("oracle-A", name, "arrivals max", arrivals.max())
    # more of the same for the cases that stress the vertex stage (tests/test_oracle_a_cpu.py has the generators):
    # an orbit view on a ragged viewport, 20:1 needles seen along an axis, a camera inside the cloud (near-face clipping)
    import math
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import test_oracle_a_cpu as T
    small, _ = scene.make_instances(8000, 0x5EED0001)
    cases = {
        "orbit": (small, O.app_camera(333, 217, eye=(5.0 * math.sin(2.0 * math.pi * 3 / 8), 2.0, 5.0 * math.cos(2.0 * math.pi * 3 / 8))), 333, 217),
        "needles": (T.anisotropic_scene(), O.app_camera(480, 270, eye=(0.0, 0.0, 6.0)), 480, 270),
        "near": (T.near_plane_scene(), O.app_camera(400, 300, eye=(0.05, 0.02, 0.3), target=(0.0, 0.0, -1.0)), 400, 300),
    }
    for cname, (ci, (cv, cp), cw, ch) in cases.items():
        out[f"ora2_{cname}_instances"] = ci; out[f"ora2_{cname}_view"] = cv; out[f"ora2_{cname}_proj"] = cp
        out[f"ora2_{cname}_wh"] = np.array([cw, ch])
        for name, mode in (("tile", 0), ("layer50", 1)):
            img, arrivals = A.render(ci, cv, cp, cw, ch, mode)
            out[f"ora2_{cname}_{name}_image"] = img
            out[f"ora2_{cname}_{name}_arrivals"] = arrivals.astype(np.uint8 if arrivals.max() < 256 else np.uint16)
            print("oracle-A", cname, name, "arrivals max", arrivals.max())
else:
    raise SystemExit("oracle/_ref/libref_metal.so missing: run `make -C oracle` where /root/reference exists")
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden_small.npz"), **out)
