#!/usr/bin/env python
"""Split the SASS-level source page of an .ncu-rep into regions delimited by BAR
instructions and report instructions executed / stall samples per region."""
import csv, io, subprocess, sys
rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
lines = out.splitlines()
start = next(i for i, l in enumerate(lines) if l.startswith('"Address"'))
rows = list(csv.DictReader(io.StringIO("\n".join(lines[start:]))))
tot_i = sum(int(r["Instructions Executed"]) for r in rows); tot_s = sum(int(r["# Samples"]) for r in rows)
print("total warp-instr", tot_i, "samples", tot_s)
reg = []; cur = dict(n=0, inst=0, samp=0, first=None, ops={}, stalls={})
stall_cols = [c for c in rows[0].keys() if c.startswith("stall_") and "Not Issued" not in c]
for r in rows:
    src = r["Source"].strip(); op = src.split()[0] if not src.startswith("@") else src.split()[1]
    cur["n"] += 1; cur["inst"] += int(r["Instructions Executed"]); cur["samp"] += int(r["# Samples"])
    base = op.split(".")[0]
    cur["ops"][base] = cur["ops"].get(base, 0) + int(r["Instructions Executed"])
    for c in stall_cols:
        cur["stalls"][c] = cur["stalls"].get(c, 0) + int(r[c] or 0)
    if cur["first"] is None: cur["first"] = src
    if base in ("BAR", "EXIT") :
        cur["last"] = src; reg.append(cur); cur = dict(n=0, inst=0, samp=0, first=None, ops={}, stalls={})
if cur["n"]: cur["last"] = "end"; reg.append(cur)
for i, g in enumerate(reg):
    if g["inst"] < 0.002 * tot_i and g["samp"] < 0.002 * tot_s: continue
    ops = sorted(g["ops"].items(), key=lambda kv: -kv[1])[:8]
    st = sorted(g["stalls"].items(), key=lambda kv: -kv[1])[:4]
    print(f"region {i:3d}: sass {g['n']:5d} inst {100*g['inst']/tot_i:5.1f}% samples {100*g['samp']/tot_s:5.1f}%  ends [{g['last'][:40]}]")
    print("     ops:", " ".join(f"{k}:{100*v/tot_i:.1f}%" for k, v in ops))
    print("     stalls:", " ".join(f"{k[6:]}:{100*v/tot_s:.1f}%" for k, v in st))
