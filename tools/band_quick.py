"""Tuning aid: stage times of the whole c3 frame and of two of its eight even bands on one GPU."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from gaussian_splat_b200 import camera, capi, scene
n, w, h = 6_000_000, 3840, 2160
inst, _ = scene.make_instances(n, scene.CONFIG_SEEDS[3])
ctx = capi.Context(0); ctx.upload(inst)
s = torch.cuda.current_stream().cuda_stream
fb = torch.zeros((h, w, 4), dtype=torch.float16, device="cuda")
V, P = camera.app_camera(w, h)
for (a, b) in [(0, 135), (34, 55), (72, 86)]:
    o = capi.make_opts(tile_row_begin=a, tile_row_end=b, flags=capi.FLAG_TIMING)
    acc = {}
    for it in range(25):
        ctx.render(s, fb.data_ptr(), V, P, w, h, o); torch.cuda.synchronize()
        if it == 0: ctx.status()
        if it >= 5:
            for k, v in ctx.stage_times().items():
                acc[k] = acc.get(k, 0.0) + v / 20
    st = ctx.status()
    print(json.dumps({"rows": [a, b], "pairs": int(st.n_pairs), "live": int(st.n_live_blocks), **{k: round(v * 1e3, 1) for k, v in acc.items()}}))
