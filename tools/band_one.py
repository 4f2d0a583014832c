"""Renders ONE tile-row band of the 6M / 4K frame a few times (ncu target for the band-mode kernels)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gaussian_splat_b200 import camera, capi, scene
a, b = int(sys.argv[1]), int(sys.argv[2])
n, w, h = 6_000_000, 3840, 2160
inst, _ = scene.make_instances(n, scene.CONFIG_SEEDS[3])
ctx = capi.Context(0); ctx.upload(inst)
s = torch.cuda.current_stream().cuda_stream
fb = torch.zeros((h, w, 4), dtype=torch.float16, device="cuda")
V, P = camera.app_camera(w, h)
for _ in range(3):
    ctx.render(s, fb.data_ptr(), V, P, w, h, capi.make_opts(tile_row_begin=a, tile_row_end=b)); torch.cuda.synchronize(); ctx.status()
print("pairs", ctx.status().n_pairs)
