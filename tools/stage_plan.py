#!/usr/bin/env python3
"""tools/stage_plan.py <key> <path.jsfx> [cut_var ...] -- BUILD-TIME helper for za_gpu/jsfxc.py::STAGE_CUTS: without cut variables,
list the top-level statements of @sample (index, emitted lines, what they assign); with them, run the planner and print each
stage's size and what it imports, or why the cuts are not legal (+key: allow a special key in several stages; ^name: hoist
the last top-level assignment to name)."""
import sys
from pathlib import Path
REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO / "zorakaudio-experimental-plugins_b200"))
from za_gpu import jsfxc  # noqa: E402

key, path, cuts = sys.argv[1], sys.argv[2], sys.argv[3:]
u = jsfxc.Unit(key, Path(path).read_text(errors="replace"))
prog, sl = u.progs["sample"], u.stmt_lines["sample"] + [len(u.body["sample"])]
names = {v: k for k, v in u.vars.items()}
nm = lambda k: names[int(k[2:-1])] if k.startswith("v[") else k
if not cuts:
    for i, n in enumerate(prog):
        what = n[2][1] if n[0] == "assign" and n[2][0] == "var" else n[0]
        print(f"{i:4d} {sl[i + 1] - sl[i]:5d} lines  {what}")
    sys.exit(0)
allow = tuple(c[1:] for c in cuts if c.startswith("+"))
moves = tuple(c[1:] for c in cuts if c.startswith("^"))          # ^name: hoist the last assignment to name
cuts = [c for c in cuts if c[0] not in "+^"]
try:
    nch = max(u.n_in, u.n_out, (max(u.used_spl) + 1) if u.used_spl else 0, 1)
    p = jsfxc.plan_stages(u, cuts, nch, u.n_out, allow, moves)
except ValueError as e:
    sys.exit(str(e))
b = [0] + p.cuts + [len(prog)]
print(f"K = {p.K}, ring slots {p.slots}, special {p.special}")
for k in range(p.K):
    print(f"stage {k}: statements {b[k]}..{b[k + 1] - 1}, {sum(sl[i + 1] - sl[i] for i in p.order[b[k]:b[k + 1]])} emitted lines, imports {[(nm(a), d) for a, o, d in p.imports[k]]}")
