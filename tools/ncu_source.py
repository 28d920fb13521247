#!/usr/bin/env python3
"""tools/ncu_source.py <report.ncu-rep> [N] -- per-opcode and top-N per-instruction warp-stall samples
from the source page of an ncu capture (needs -lineinfo / --import-source on)."""
import collections, csv, subprocess, sys
rep, topn = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 15
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hi]
col = {c: i for i, c in enumerate(h)}
stalls = [c for c in h if c.startswith("stall_") and "Not Issued" not in c]
data = []
for r in rows[hi + 1:]:
    if len(r) < len(h) or not r[0].startswith("0x"):
        continue
    data.append(r)
tot = sum(int(r[col["# Samples"]]) for r in data)
print("instructions:", len(data), " total samples:", tot)
agg = collections.Counter(); st = collections.Counter()
for r in data:
    toks = r[col["Source"]].split()
    op = toks[1] if toks and toks[0].startswith("@") else (toks[0] if toks else "")
    agg[op.split(".")[0]] += int(r[col["# Samples"]])
    for s in stalls:
        st[s] += int(r[col[s]] or 0)
print("by stall reason:", ", ".join(f"{k[6:]} {100 * v / tot:.1f}%" for k, v in st.most_common(9)))
print("by opcode:", ", ".join(f"{k} {100 * v / tot:.1f}%" for k, v in agg.most_common(12)))
for r in sorted(data, key=lambda r: -int(r[col["# Samples"]]))[:topn]:
    top = sorted(((int(r[col[s]] or 0), s[6:]) for s in stalls), reverse=True)[:2]
    print(f"{int(r[col['# Samples']]):6d}  {r[col['Source']].strip():70s} {top}")
