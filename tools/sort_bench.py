"""Stand-alone sort benchmark (the 'sort HBM GB/s' metric): python tools/sort_bench.py [n] [end_bit] [reps] [key_bits]
key_bits = 64 (default) or 32 (the width frames sort: 16 B per pair and pass instead of 24)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from gaussian_splat_b200 import capi

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 26
end_bit = int(sys.argv[2]) if len(sys.argv) > 2 else 64
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
key_bits = int(sys.argv[4]) if len(sys.argv) > 4 else 64
end_bit = min(end_bit, key_bits)
ctx = capi.Context(0)
g = torch.Generator(device="cuda"); g.manual_seed(1)
hi = torch.randint(-2 ** 31, 2 ** 31, (n,), generator=g, dtype=torch.int64, device="cuda")
lo = torch.randint(0, 2 ** 32, (n,), generator=g, dtype=torch.int64, device="cuda")
src = (hi << 32) | lo
if end_bit < 64:
    src &= (1 << end_bit) - 1
if key_bits == 32:
    src = src.to(torch.int32)      # the low 32 bits
del hi, lo
k = torch.empty_like(src); ka = torch.empty_like(src)
v = torch.arange(n, dtype=torch.int32, device="cuda"); va = torch.empty_like(v)
s = torch.cuda.current_stream().cuda_stream
ts = []
for it in range(reps + 2):
    k.copy_(src)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    alt = ctx.sort_pairs(s, k.data_ptr(), v.data_ptr(), ka.data_ptr(), va.data_ptr(), n, 0, end_bit, key_bits=key_bits)
    e1.record(); torch.cuda.synchronize()
    if it >= 2:
        ts.append(e0.elapsed_time(e1))
out = ka if alt else k
assert bool((out[1:] >= out[:-1]).all()) if end_bit < key_bits else True
n_pass = (end_bit + 7) // 8
ms = float(np.median(ts)); by = n * (key_bits // 8) * (1 + n_pass * 2) + n * n_pass * 8
print(f"key_bits={key_bits} n={n} passes={n_pass} ms={ms:.4f} us/pass={1e3*ms/n_pass:.1f} GB/s={by/ms/1e6:.1f} Mkeys/s={n/ms/1e3:.1f}")
