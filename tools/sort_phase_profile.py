"""Per-phase clock64 breakdown of one onesweep pass (needs the -DSORT_PROFILE build: libsplat_cuda_prof.so)."""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from gaussian_splat_b200 import capi
capi.LIB_PATH = os.path.join(os.path.dirname(capi.LIB_PATH), "libsplat_cuda_prof.so")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 26
key_bits = int(sys.argv[2]) if len(sys.argv) > 2 else 64
tile_keys = int(sys.argv[3]) if len(sys.argv) > 3 else (4096 if key_bits == 64 else 6144)
ctx = capi.Context(0)
g = torch.Generator(device="cuda"); g.manual_seed(1)
src = torch.randint(0, 2 ** 62, (n,), generator=g, dtype=torch.int64, device="cuda")
if key_bits == 32:
    src = src.to(torch.int32)
k = src.clone(); ka = torch.empty_like(k); v = torch.arange(n, dtype=torch.int32, device="cuda"); va = torch.empty_like(v)
s = torch.cuda.current_stream().cuda_stream
for it in range(3):
    k.copy_(src)
    ctx.sort_pairs(s, k.data_ptr(), v.data_ptr(), ka.data_ptr(), va.data_ptr(), n, 0, 8, key_bits=key_bits)   # ONE pass
    torch.cuda.synchronize()
h = ctx.fetch(capi.FETCH_DIGIT_HISTOGRAMS)[7, :8].astype(np.float64) * 16
names = ["ticket", "load", "rank", "scan", "scatter_smem", "lookback", "writeout+loop_sync", "-"]
tot = h[:7].sum()
for nm, c in zip(names, h[:7]):
    print(f"{nm:20s} {c / tot * 100:5.1f}%   {c / max(1, (n + tile_keys - 1) // tile_keys):9.0f} cycles/tile")
