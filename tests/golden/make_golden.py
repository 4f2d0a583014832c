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
else:
    raise SystemExit("oracle/_ref/libref_metal.so missing: run `make -C oracle` where /root/reference exists")
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden_small.npz"), **out)
