"""Per-instruction stall samples of one kernel from an ncu report: python tools/ncu_source.py X.ncu-rep <kernel regex> [min samples]"""
import csv, subprocess, sys
rep, pat = sys.argv[1], sys.argv[2]
mins = int(sys.argv[3]) if len(sys.argv) > 3 else 20
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "-k", "regex:" + pat], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr = None; tot = 0; out = []
for r in rows:
    if r and r[0] == "Address": hdr = r; continue
    if r and r[0] == "Kernel Name": out.append(("K", r[1][:80])); continue
    if hdr is None or len(r) < 6: continue
    try: s = int(r[2]); ex = int(r[5])
    except ValueError: continue
    tot += s
    out.append((s, ex, r[1].strip()[:110]))
print("total samples", tot)
for o in out:
    if o[0] == "K": print("==", o[1])
    elif o[0] >= mins: print("%6d %9d  %s" % o)
