#!/usr/bin/env python3
"""tools/prof_run.py -- short device-resident run of one plugin for ncu / timing sweeps.
   python tools/prof_run.py --plugin cascade8 --streams 1024 --seconds 10 --precision f32 --reps 3 [--warps W]
Prints per-rep kernel ms (CUDA events inside the library) and algorithmic GB/s."""
import argparse, json, sys
from pathlib import Path
REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO / "zorakaudio-experimental-plugins_b200"))
import numpy as np
import torch
import za_gpu
from za_gpu import workloads as wl

ap = argparse.ArgumentParser()
ap.add_argument("--plugin", default="cascade8")
ap.add_argument("--streams", type=int, default=1024)
ap.add_argument("--seconds", type=float, default=10.0)
ap.add_argument("--precision", default="f32")
ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--warps", type=int, default=0)
ap.add_argument("--opt", action="append", default=[], help="name=value library option")
a = ap.parse_args()
info = za_gpu.plugin_info(a.plugin)
ns, nf = a.streams, int(a.seconds * 48000)
x = (torch.rand((ns, info.n_in, nf), device="cuda") - 0.5) * 0.5
y = torch.empty((ns, info.n_out, nf), device="cuda")
params = wl.default_params(a.plugin, ns)
bp = za_gpu.BatchProcessor(a.plugin, ns, 48000.0, 512, a.precision, 0)
if a.warps:
    bp.set_option("warps_per_stream", a.warps)
for o in a.opt:
    k, v = o.split("=")
    bp.set_option(k, int(v))
bp.prepare(params)
bpf = 4 * (info.n_in + info.n_out)
res = []
for r in range(a.reps):
    bp.process(x, out=y)
    ms, n = bp.last_kernel_ms()
    res.append(ms)
    print(f"rep {r}: kernel {ms:.3f} ms over {n} launch(es): {ns * nf * bpf / ms / 1e6:.1f} GB/s algorithmic", flush=True)
print(json.dumps({"plugin": a.plugin, "precision": a.precision, "streams": ns, "frames": nf, "warps": a.warps,
                  "best_ms": min(res), "gbps": ns * nf * bpf / min(res) / 1e6}))
bp.close()
