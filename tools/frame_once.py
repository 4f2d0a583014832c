"""Renders one bench workload's frame a few times (ncu --set full target: every kernel of the frame once)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from gaussian_splat_b200 import camera, capi, scene
wl = sys.argv[1] if len(sys.argv) > 1 else "c2"
n, w, h, mode, cfg, name = bench.WORKLOADS[wl]
inst, _ = scene.make_instances(n, scene.CONFIG_SEEDS[cfg])
ctx = capi.Context(0); ctx.upload(inst)
s = torch.cuda.current_stream().cuda_stream
fb = torch.zeros((h, w, 4), dtype=torch.float16, device="cuda")
V, P = camera.app_camera(w, h)
for _ in range(4):
    ctx.render(s, fb.data_ptr(), V, P, w, h, capi.make_opts(mode=mode)); torch.cuda.synchronize(); ctx.status()
print(name, "pairs", ctx.status().n_pairs)
