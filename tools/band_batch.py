"""Tuning aid: one band of eight of the 6M / 4K frame, frames back to back: splat_render per frame on one stream against
splat_render_batch (two frames in flight on two frame sets when the library's rule says so; SPLAT_TWO_FRAMES=0/1 forces)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from gaussian_splat_b200 import camera, capi, scene
n, w, h = 6_000_000, 3840, 2160
inst, _ = scene.make_instances(n, scene.CONFIG_SEEDS[3])
ctx = capi.Context(0); ctx.upload(inst)
s = torch.cuda.current_stream().cuda_stream
fbs = [torch.zeros((h, w, 4), dtype=torch.float16, device="cuda") for _ in range(4)]
V, P = camera.app_camera(w, h)
K = 48
for (a, b) in [(34, 55), (72, 86), (0, 135)]:
    o = capi.make_opts(tile_row_begin=a, tile_row_end=b)
    for _ in range(3):
        ctx.render(s, fbs[0].data_ptr(), V, P, w, h, o); torch.cuda.synchronize(); ctx.status()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(K):
        ctx.render(s, fbs[i & 3].data_ptr(), V, P, w, h, o)
    e1.record(); torch.cuda.synchronize()
    serial = e0.elapsed_time(e1) / K * 1e3
    views = np.stack([V] * K); projs = np.stack([P] * K)
    ptrs = [fbs[i & 3].data_ptr() for i in range(K)]
    ctx.render_batch(s, ptrs[:4], views[:4], projs[:4], w, h, o); torch.cuda.synchronize()
    e0.record()
    ctx.render_batch(s, ptrs, views, projs, w, h, o)
    e1.record(); torch.cuda.synchronize()
    batch = e0.elapsed_time(e1) / K * 1e3
    ref = fbs[0].clone()
    ctx.render(s, fbs[1].data_ptr(), V, P, w, h, o); torch.cuda.synchronize()
    a0, a1 = a * 16, min(b * 16, h)
    print(json.dumps({"rows": [a, b], "serial_us": round(serial, 1), "batch_us": round(batch, 1), "equal": bool(torch.equal(ref[a0:a1], fbs[1][a0:a1]))}))
