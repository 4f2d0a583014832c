"""SASS opcode histogram per kernel of the built objects (no GPU needed): python tools/sass_histogram.py > profiles/rNN_sass_opcodes.txt
Shows what the evidence in DESIGN.md claims: PREEXIT / ACQBULK (programmatic dependent launch), UBLKPF (TMA bulk L2 prefetch in
the sort), VOTE / SHFL / MATCH, ATOMS / ATOMG / RED, the half-precision ops of the blend, and the absence of tensor-core
(HMMA / UTC*MMA) and TMA-load (UTMALDG) instructions on a path that has no dense contraction."""
import collections, glob, os, re, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WATCH = ["PREEXIT", "ACQBULK", "UBLKPF", "UTMALDG", "HMMA", "UTCHMMA", "MATCH", "VOTE", "SHFL", "ATOMS", "ATOMG", "RED", "LDG", "STG", "LDS", "STS",
         "BAR", "MUFU", "FFMA", "FMUL", "FADD", "HFMA2", "HMUL2", "HADD2", "F2F", "F2FP", "LDL", "STL"]
for obj in sorted(glob.glob(os.path.join(ROOT, "gaussian_splat_b200", "csrc", "build", "*.o"))):
    txt = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
    fn, hist = None, collections.OrderedDict()
    for ln in txt.splitlines():
        m = re.match(r"\s*Function : (\S+)", ln)
        if m:
            fn = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
            hist[fn] = collections.Counter(); continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", ln)
        if m and fn:
            hist[fn][m.group(1)] += 1
    if not hist:
        continue
    print("==", os.path.basename(obj))
    for fn, c in hist.items():
        total = sum(c.values())
        sel = " ".join(f"{k}={c[k]}" for k in WATCH if c.get(k))
        print(f"  {fn[:90]:<90} total={total} {sel}")
