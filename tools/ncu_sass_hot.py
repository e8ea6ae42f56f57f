#!/usr/bin/env python
"""Where does a kernel's time go, from the SASS source page of an .ncu-rep
(`ncu -i rep --page source --csv --print-source sass --kernel-name regex:K`):
samples per block of BLOCK instructions with the stall-reason split and the opcode mix,
then the hottest single instructions.  Usage: tools/ncu_sass_hot.py rep kernel_regex [block]"""
import csv
import io
import subprocess
import sys
from collections import Counter


def main():
    rep, kern = sys.argv[1], sys.argv[2]
    block = int(sys.argv[3]) if len(sys.argv) > 3 else 100
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "--kernel-name",
                          "regex:" + kern], capture_output=True, text=True).stdout
    lines = out.splitlines()
    start = next(i for i, l in enumerate(lines) if l.startswith('"Address"'))
    rows = [r for r in csv.DictReader(io.StringIO("\n".join(lines[start:]))) if (r.get("# Samples") or "").isdigit()]
    stalls = [k for k in rows[0] if k.startswith("stall_") and "Not Issued" not in k]
    tot = sum(int(r["# Samples"]) for r in rows)
    print(f"# {len(rows)} SASS instructions, {tot} samples")
    for b0 in range(0, len(rows), block):
        rs = rows[b0:b0 + block]
        s = sum(int(r["# Samples"]) for r in rs)
        if s < tot * 0.004:
            continue
        st = Counter()
        for r in rs:
            for k in stalls:
                st[k[6:]] += int(r[k] or 0)
        ops = Counter(r["Source"].split()[0].lstrip("@!P0123456789T ").split(".")[0] if not r["Source"].strip().startswith("@")
                      else r["Source"].split()[1].split(".")[0] for r in rs)
        ex = sum(int(r["Instructions Executed"]) for r in rs)
        top = ", ".join(f"{k} {100 * v / max(1, sum(st.values())):.0f}%" for k, v in st.most_common(4))
        mix = " ".join(f"{k}:{v}" for k, v in ops.most_common(6))
        print(f"[{b0:5d}-{b0 + len(rs):5d}) {100 * s / tot:5.1f}%  exec {ex / 1e6:8.1f}M  | {top} | {mix}")
    print("# hottest instructions")
    for i, r in sorted(enumerate(rows), key=lambda ir: -int(ir[1]["# Samples"]))[:25]:
        st = Counter({k[6:]: int(r[k] or 0) for k in stalls})
        top = ", ".join(f"{k} {v}" for k, v in st.most_common(2))
        print(f"{i:5d} {100 * int(r['# Samples']) / tot:5.2f}%  {r['Source'].strip()[:70]:70s} {top}")


if __name__ == "__main__":
    main()
