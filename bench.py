#!/usr/bin/env python3
"""bench.py -- BASELINE.json metric on the B200 batch renderer.

Workload at N=1 (BASELINE config 2, `configs[1]`): plugin `cascade8` (8-section biquad/one-pole EQ
cascade), 1024 mono streams x 60 s at 48 kHz = 2.949e9 sample-frames per step, float32 planar I/O,
per-stream seeded parameters.  One step = one za_gpu_process() pass over the whole batch (state
carries from step to step, as in a continuing render).  N>1: every rank renders its own 1024
streams on its own GPU (streams shard, no collective on the data path) -> weak scaling.

  value     frames/s with inputs resident in HBM (CUDA events around K back-to-back steps)
  e2e       the same metric through the C ABI with HOST buffers: pinned input -> H2D -> kernels ->
            D2H -> pinned output inside the timed region (za_gpu_process(..., ZA_MEM_HOST))
  roofline  algorithmic bytes (8 B per mono frame: 4 in + 4 out) / dominant-kernel time measured
            with CUDA events inside the timed region, against MEASURED_PEAKS.json hbm_gbs
  cpu_baseline  the reference's own compiled DSP-JSFX object (oracle/_ref, else the pinned C port)
            multithreaded over a bounded sample of the same workload on this box's host cores

`--impl reference` times that CPU path alone (rank 0) on the same metric/config.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

REPO = Path(__file__).resolve().parent
sys.path.insert(0, str(REPO / "zorakaudio-experimental-plugins_b200"))

SRATE = 48000.0
N_STREAMS = 1024
SECONDS = 60
BYTES_PER_FRAME = 8           # mono float32 in + out
CPU_SAMPLE_STREAMS = 1024     # bounded CPU sample: all 1024 streams x 20 s (a third of the workload)
CPU_SAMPLE_SECONDS = 20.0


def peaks():
    p = REPO / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.rows, self.proc, self.idx = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.idx)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm = [float(r[1]) for r in self.rows if len(r) >= 8 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 8 and r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 8 for n, v in zip(names, r[4:8]) if v.lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def cpu_reference_run(n_streams: int, seconds: float, threads: int):
    """Reference CPU path on a bounded sample of the workload -> (frames/s, kind, seconds taken)."""
    sys.path.insert(0, str(REPO / "oracle"))
    import za_oracle as zo
    from za_gpu import workloads as wl
    orc = zo.best_oracle("cascade8")
    kind = "reference" if isinstance(orc, zo.RefPlugin) else "port"
    nf = int(seconds * SRATE)
    p, x = wl.cascade8_params(n_streams), wl.noise_two_sines(min(n_streams, 8), nf)
    x = np.ascontiguousarray(np.tile(x, ((n_streams + x.shape[0] - 1) // x.shape[0], 1, 1))[:n_streams])
    orc.render(p[:threads], x[:threads, :, :4800], threads=threads)          # warm
    t0 = time.perf_counter()
    orc.render(p, x, block=512, threads=threads)
    dt = time.perf_counter() - t0
    return n_streams * nf / dt, kind, dt


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    ns, secs = CPU_SAMPLE_STREAMS, CPU_SAMPLE_SECONDS
    vals = []
    for i in range(args.warmup + args.steps):
        v, kind, dt = cpu_reference_run(ns, secs, threads)
        if i >= args.warmup:
            vals.append((v, dt))
    v = float(np.mean([a for a, _ in vals]))
    sample = f"{ns} streams x {secs:.0f} s per step (of {N_STREAMS} x {SECONDS} s), block 512, FTZ/DAZ on"
    line = {
        "impl": "reference", "metric": "aggregate sample-frames/s", "value": v, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": float(np.mean([d for _, d in vals]) * 1e3),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "x_realtime": v / SRATE,
        "config": {"workload": "cfg2: cascade8 (8-section biquad/one-pole cascade), 1024 mono streams x 60 s @ 48 kHz",
                   "sample": sample},
        "cpu_baseline": {"value": v, "unit": "frames/s", "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--precision", default="f32", choices=["f32", "f64"])
    ap.add_argument("--streams", type=int, default=N_STREAMS)
    ap.add_argument("--seconds", type=float, default=SECONDS)
    ap.add_argument("--e2e-streams", type=int, default=0, help="0 = the full batch")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--warps", type=int, default=0)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    if args.impl == "reference":
        run_reference_arm(args)
        return

    import torch
    import za_gpu

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the renderer has no CPU fallback")
    torch.cuda.set_device(local)
    # a real (non-NULL) stream: the library enqueues on it and torch's events time it
    run_stream = torch.cuda.Stream()
    torch.cuda.set_stream(run_stream)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    ns, nf = args.streams, int(args.seconds * SRATE)
    from za_gpu import workloads as wl
    params = wl.cascade8_params(ns, first_stream=rank * ns)

    # ---- synthetic input, generated on the device: uniform noise x0.25 + two sines per stream -------
    g = torch.Generator(device="cuda").manual_seed(5000 + rank)
    x = torch.empty((ns, 1, nf), dtype=torch.float32, device="cuda")
    t = torch.arange(nf, device="cuda", dtype=torch.float32) / SRATE
    f = torch.rand((ns, 2), device="cuda", generator=g)
    for s0 in range(0, ns, 64):
        s1 = min(ns, s0 + 64)
        blk = (torch.rand((s1 - s0, 1, nf), device="cuda", generator=g) - 0.5) * 0.5
        f1 = (60.0 + 340.0 * f[s0:s1, 0]).view(-1, 1, 1)
        f2 = (1000.0 + 5000.0 * f[s0:s1, 1]).view(-1, 1, 1)
        blk += 0.125 * torch.sin(2 * np.pi * f1 * t) + 0.125 * torch.sin(2 * np.pi * f2 * t)
        x[s0:s1] = blk
    y = torch.empty_like(x)
    del blk
    torch.cuda.synchronize()

    results = {}
    for prec in ([args.precision] + (["f64"] if args.precision == "f32" else [])):
        bp = za_gpu.BatchProcessor("cascade8", ns, SRATE, 512, prec, local)
        if args.warps:
            bp.set_option("warps_per_stream", args.warps)
        bp.set_stream(torch.cuda.current_stream().cuda_stream)
        bp.prepare(params)
        headline = prec == args.precision
        steps = args.steps if headline else max(2, args.steps // 2)
        # parity spot check on the very first pass (fresh state): 4 streams x 1 s against the oracle
        bp.process(x, out=y)
        torch.cuda.synchronize()
        parity = None
        if rank == 0:
            try:
                sys.path.insert(0, str(REPO / "oracle"))
                import za_oracle as zo
                pick = [0, ns // 3, ns // 2, ns - 1]
                xs = x[pick, :, :48000].cpu().numpy()
                yr, yrd = zo.best_oracle("cascade8").render(params[pick], xs, want_f64=True)
                got = y[pick, :, :48000].cpu().numpy()
                r = got.astype(np.float64) - yrd
                parity = {"streams": pick, "frames": 48000, "float_equal_frac": float(np.mean(got == yr)),
                          "residual_dbfs_rms": float(20 * np.log10(np.sqrt(np.mean(r ** 2)) + 1e-300))}
            except Exception as e:      # the checker is optional for the measurement
                parity = {"error": str(e)[:120]}
        for _ in range(args.warmup - 1):
            bp.process(x, out=y)
        barrier()
        clk = ClockSampler(local)
        if headline and rank == 0:
            clk.start()
        l0 = bp.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        kms, klaunch = 0.0, 0
        e0.record()
        for _ in range(steps):
            bp.process(x, out=y)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        k1, kn = bp.last_kernel_ms()           # dominant kernel of the LAST timed step (CUDA events on the stream)
        kms, klaunch = k1, kn
        launches = bp.launch_count() - l0
        clocks = clk.stop() if (headline and rank == 0) else None
        tmax = torch.tensor([ms], device="cuda")
        if dist is not None:
            dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        ms_max = float(tmax.item())
        results[prec] = dict(ms=ms_max, value=world * ns * nf / (ms_max * 1e-3), kernel_ms=kms, kernel_launches=klaunch,
                             launches=launches, clocks=clocks, parity=parity, steps=steps)
        bp.close()

    # ---- e2e: HOST buffers through the C ABI, copies inside the timed region ---------------------------
    e2e = None
    es = args.e2e_streams or ns
    while es >= 16:
        try:
            hx = torch.empty((es, 1, nf), dtype=torch.float32, pin_memory=True)
            hy = torch.empty((es, 1, nf), dtype=torch.float32, pin_memory=True)
            break
        except RuntimeError:
            es //= 4
            hx = hy = None
    if hx is not None:
        hx.copy_(x[:es])
        torch.cuda.synchronize()
        bp = za_gpu.BatchProcessor("cascade8", es, SRATE, 512, args.precision, local).prepare(params[:es])
        for _ in range(2):
            bp.process(hx, out=hy)
        barrier()
        t0 = time.perf_counter()
        e_steps = max(2, args.steps // 2)
        for _ in range(e_steps):
            bp.process(hx, out=hy)              # returns when `hy` is complete
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / e_steps
        tm = torch.tensor([dt], device="cuda")
        if dist is not None:
            dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        dt = float(tm.item())
        same = bool(torch.equal(hy[:4].cuda(), y[:4])) if args.precision in results and False else None
        e2e = {"value": world * es * nf / dt, "unit": "frames/s", "h2d_bytes_per_step": es * nf * 4,
               "d2h_bytes_per_step": es * nf * 4, "ms_per_step": dt * 1e3, "streams_per_gpu": es,
               "x_realtime": world * es * nf / dt / SRATE, "host_memory": "pinned"}
        bp.close()
        del hx, hy

    if rank != 0:
        if dist is not None:
            dist.barrier()
            dist.destroy_process_group()
        return

    peak, peak_kind = peaks()
    r = results[args.precision]
    traffic = None
    tp = REPO / "profiles" / "cascade8_traffic.json"
    if tp.exists():
        try:
            traffic = json.loads(tp.read_text()).get(args.precision, {}).get("dram_bytes_per_launch")
        except Exception:
            traffic = None

    def roof(rr):
        if not rr["kernel_launches"] or rr["kernel_ms"] <= 0:
            return None
        per_launch_ms = rr["kernel_ms"] / rr["kernel_launches"]
        ach = ns * nf * BYTES_PER_FRAME / rr["kernel_launches"] * rr["kernel_launches"] / (rr["kernel_ms"] * 1e-3) / 1e9
        return {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "peak_source": peak_kind,
                "traffic": traffic, "kernel": "cascade8_kernel", "kernel_ms_per_launch": per_launch_ms,
                "algorithmic_bytes_per_launch": ns * nf * BYTES_PER_FRAME // max(rr["kernel_launches"], 1)}

    cpu = None
    if not args.no_cpu:
        threads = os.cpu_count() or 1
        cns = CPU_SAMPLE_STREAMS
        v, kind, dt = cpu_reference_run(cns, CPU_SAMPLE_SECONDS, threads)
        cpu = {"value": v, "unit": "frames/s", "cores": threads, "kind": kind,
               "sample": f"{cns} streams x {CPU_SAMPLE_SECONDS:g} s of the same workload ({dt:.1f} s of wall time), block 512",
               "x_realtime": v / SRATE}

    line = {
        "metric": "aggregate sample-frames/s", "value": r["value"], "unit": "frames/s", "n_gpus": world,
        "steps": r["steps"], "warmup": args.warmup, "ms_per_step": r["ms"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": args.precision, "data": "synthetic",
        "x_realtime": r["value"] / SRATE,
        "config": {"workload": f"cfg2: cascade8 (8-section biquad/one-pole cascade), {ns} mono streams x {args.seconds:g} s @ 48 kHz per GPU",
                   "streams_per_gpu": ns, "frames_per_stream": nf, "block_len": 512, "io": "float32 planar [stream][ch][frame]",
                   "l2": "inputs (%.1f GB per step) larger than L2" % (ns * nf * 4 / 1e9), "parallelism": f"streams sharded x{world}"},
        "clocks": r["clocks"], "e2e": e2e, "gpu_launches": r["launches"], "roofline": roof(r), "cpu_baseline": cpu,
        "parity_spot_check": r["parity"],
    }
    if "f64" in results and args.precision != "f64":
        q = results["f64"]
        line["f64"] = {"value": q["value"], "ms_per_step": q["ms"], "x_realtime": q["value"] / SRATE, "roofline": roof(q),
                       "parity_spot_check": q["parity"], "gpu_launches": q["launches"]}
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
