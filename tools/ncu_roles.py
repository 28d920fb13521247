#!/usr/bin/env python3
"""tools/ncu_roles.py <report.ncu-rep> -- warp-stall samples of a warp-specialised kernel per code segment: the SASS
is cut at every BAR.SYNC (each role's loop ends in one), so a segment = one role's loop body."""
import collections, csv, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hi]; col = {c: i for i, c in enumerate(h)}
stalls = [c for c in h if c.startswith("stall_") and "Not Issued" not in c]
segs, cur = [], {"n": 0, "samples": 0, "st": collections.Counter(), "ops": collections.Counter(), "exec": 0}
ex = col.get("Warp Instructions Executed") or col.get("Instructions Executed")
for r in rows[hi + 1:]:
    if len(r) < len(h) or not r[0].startswith("0x"):
        continue
    src = r[col["Source"]]
    cur["n"] += 1; cur["samples"] += int(r[col["# Samples"]])
    if ex is not None:
        cur["exec"] += int(r[ex] or 0)
    toks = src.split(); op = toks[1] if toks and toks[0].startswith("@") else (toks[0] if toks else "")
    cur["ops"][op.split(".")[0]] += int(r[col["# Samples"]])
    for s in stalls:
        cur["st"][s[6:]] += int(r[col[s]] or 0)
    if "BAR.SYNC" in src or "EXIT" in src:
        segs.append(cur); cur = {"n": 0, "samples": 0, "st": collections.Counter(), "ops": collections.Counter(), "exec": 0}
segs.append(cur)
tot = sum(s["samples"] for s in segs)
for i, s in enumerate(segs):
    if s["samples"] < tot * 0.002:
        continue
    print(f"seg {i:2d}: {s['n']:5d} instr, warp-instr executed {s['exec']:>12d}, samples {100 * s['samples'] / tot:5.1f}%  " +
          ", ".join(f"{k} {100 * v / max(s['samples'], 1):.0f}%" for k, v in s["st"].most_common(5)))
