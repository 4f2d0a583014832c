"""Per-stage times of every tile-row band of a G-way split of the 6M / 4K frame, measured on ONE GPU
(what each rank of `bench.py --workload c3 --gpus G` executes per frame, without the gather)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from gaussian_splat_b200 import camera, capi, multigpu, scene

G = int(sys.argv[1]) if len(sys.argv) > 1 else 8
n, w, h = 6_000_000, 3840, 2160
if len(sys.argv) > 2:
    n = int(sys.argv[2])
inst, _ = scene.make_instances(n, scene.CONFIG_SEEDS[3])
ctx = capi.Context(0); ctx.upload(inst)
s = torch.cuda.current_stream().cuda_stream
fb = torch.zeros((h, w, 4), dtype=torch.float16, device="cuda")
V, P = camera.app_camera(w, h)
tiles_y = (h + 15) // 16
for _ in range(2):
    ctx.render(s, fb.data_ptr(), V, P, w, h, capi.make_opts()); torch.cuda.synchronize(); ctx.status()
r = ctx.fetch(capi.FETCH_TILE_RANGES).astype(np.int64)
row_pairs = (r[:, 1] - r[:, 0]).reshape(tiles_y, -1).sum(axis=1)
bands = multigpu.band_rows(tiles_y, G, row_pairs)
out = []
for (a, b) in [(0, tiles_y)] + list(bands):
    o = capi.make_opts(tile_row_begin=a, tile_row_end=b, flags=capi.FLAG_TIMING)
    acc = {}
    for it in range(25):
        ctx.render(s, fb.data_ptr(), V, P, w, h, o); torch.cuda.synchronize()
        if it >= 5:
            for k, v in ctx.stage_times().items():
                acc[k] = acc.get(k, 0.0) + v / 20
    st = ctx.status()
    row = {"rows": [a, b], "pairs": int(st.n_pairs), "live_blocks": int(st.n_live_blocks), "slots": int(st.n_candidates), "max_list": int(st.max_tile_pairs), **{k: round(v * 1e3, 1) for k, v in acc.items()}}
    out.append(row); print(json.dumps(row))
