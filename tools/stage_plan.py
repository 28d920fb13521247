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
if cuts and cuts[0] == "--auto":             # --auto K [+key ...]: K stages with the smallest largest stage (emitted lines) over the legal cut points
    import functools
    K, allow = int(cuts[1]), tuple(c[1:] for c in cuts[2:] if c.startswith("+"))
    nch = max(u.n_in, u.n_out, (max(u.used_spl) + 1) if u.used_spl else 0, 1)
    legal = []
    for i in range(1, len(prog)):
        try:
            jsfxc.plan_stages(u, [i], nch, u.n_out, allow); legal.append(i)
        except ValueError:
            pass
    pts = [0] + legal + [len(prog)]

    @functools.lru_cache(None)
    def best(ai, k):
        if k == 1:
            return (sl[pts[-1]] - sl[pts[ai]], ())
        res = (10 ** 9, ())
        for bi in range(ai + 1, len(pts) - 1):
            c = sl[pts[bi]] - sl[pts[ai]]
            if c >= res[0]:
                break
            r = best(bi, k - 1)
            if max(c, r[0]) < res[0]:
                res = (max(c, r[0]), (pts[bi],) + r[1])
        return res
    m, found = best(0, K)
    print(f"{len(legal)} legal cut points of {len(prog) - 1}; best {K} stages: cuts {list(found)}, largest stage {m} lines")
    cuts = [str(c) for c in found] + ["+" + a_ for a_ in allow]
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
