#!/usr/bin/env python3
"""tools/launch_shares.py LABEL launches.csv [out.txt] -- per-kernel launch counts and time shares of one
`ncu --metrics gpu__time_duration.sum --clock-control none --csv` launch list (cold-cache, serialised: only the SHARE of a
kernel in the step is meaningful, not the absolute time).  Kernels of other libraries (torch fills / copies) are listed apart."""
import collections, csv, re, sys

label, path = sys.argv[1], sys.argv[2]
rows = [r for r in csv.reader(open(path, errors="replace")) if len(r) > 14 and r[0].isdigit()]
ours, other = collections.defaultdict(lambda: [0, 0.0]), collections.defaultdict(lambda: [0, 0.0])
for r in rows:
    name, ns = r[4], float(r[-1].replace(",", ""))
    short = re.sub(r"\(.*", "", name).replace("void ", "").replace("za::", "").replace("unnamed>::", "").replace("<unnamed>::", "")
    d = other if ("at::" in name or "vectorized_elementwise" in name or "cub::" in name or "nccl" in name.lower()) else ours
    d[short][0] += 1; d[short][1] += ns / 1e6
tot = sum(v[1] for v in ours.values())
lines = [f"{label}: {sum(v[0] for v in ours.values())} launches of our kernels, {tot:.2f} ms under ncu (cold-cache, serialised); "
         f"{sum(v[0] for v in other.values())} launches of other libraries' kernels (torch fills / copies of the bench harness), {sum(v[1] for v in other.values()):.2f} ms"]
for k, (n, ms) in sorted(ours.items(), key=lambda kv: -kv[1][1]):
    lines.append(f"    {k[:70]:70s} n={n:4d} {ms:11.3f} ms  share {100 * ms / max(tot, 1e-9):5.1f} %")
txt = "\n".join(lines) + "\n"
if len(sys.argv) > 3:
    open(sys.argv[3], "a").write(txt)
print(txt, end="")
