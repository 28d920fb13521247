#!/usr/bin/env python3
"""bench.py -- BASELINE.json metric on the B200 batch renderer.

Default workload at N=1 (BASELINE config 2, `configs[1]`, the one the metric is quoted on): plugin
`cascade8` (8-section biquad/one-pole EQ cascade), 1024 mono streams x 60 s at 48 kHz = 2.949e9
sample-frames per step, float32 planar I/O, per-stream seeded parameters.  One step = one
za_gpu_process() pass over the whole batch (state carries from step to step, as in a continuing
render).  N>1: every rank renders its own batch on its own GPU (streams shard, no collective on the
data path) -> weak scaling.  `--workload cfg3|cfg4` runs the same contract on the other single-GPU
configs (EasyExpander 4096 stereo x 60 s with per-block automation; stft_mask2048 512 stereo x 120 s).

  value     frames/s with inputs resident in HBM (CUDA events around K back-to-back steps)
  e2e       the same metric through the C ABI with HOST buffers: pinned input -> H2D -> kernels ->
            D2H -> pinned output inside the timed region (za_gpu_process(..., ZA_MEM_HOST))
  roofline  algorithmic bytes (4 B x (in + out channels) per frame) / dominant-kernel
            time measured with CUDA events inside the timed region, against MEASURED_PEAKS.json hbm_gbs
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

# name -> plugin, streams per GPU, seconds, channels in/out, automated slider columns, CPU sample (streams, seconds), kernel
WORKLOADS = {
    "cfg2": dict(plugin="cascade8", streams=1024, seconds=60.0, ch=(1, 1), automate=None, cpu=(1024, 20.0),
                 kernel={"f32": "cascade8_x2_kernel", "f64": "cascade8_kernel"},
                 text="cfg2: cascade8 (8-section biquad/one-pole cascade), {ns} mono streams x {sec:g} s @ 48 kHz per GPU"),
    "cfg3": dict(plugin="EasyExpander", streams=4096, seconds=60.0, ch=(2, 2), automate=[0, 1], cpu=(256, 10.0),
                 kernel={"f32": "easyexpander_kernel", "f64": "easyexpander_kernel"},
                 text="cfg3: EasyExpander (nonlinear envelope + gain computer), {ns} stereo streams x {sec:g} s @ 48 kHz per GPU, "
                      "threshold/depth automated every 512-frame block, in-place buffers"),
    "cfg4": dict(plugin="stft_mask2048", streams=512, seconds=120.0, ch=(2, 2), automate=None, cpu=(64, 10.0),
                 kernel={"f32": "stft_mask_kernel", "f64": "stft_mask_kernel"},
                 text="cfg4: stft_mask2048 (2048-pt Hann STFT, hop 512, per-bin gain, overlap-add), {ns} stereo streams x {sec:g} s @ 48 kHz per GPU"),
    # cfg3's shape on the Faust leaf of the Dynamics category (float32 only; CPU side = the restatement, parity unpinned)
    "modtilt": dict(plugin="ModTilt", streams=4096, seconds=60.0, ch=(2, 2), automate=[0], cpu=(512, 20.0), f32_only=True,
                    kernel={"f32": "modtilt_kernel"},
                    text="cfg3 shape on the Faust leaf ModTilt (envelope tilt shaper: 5 one-poles + sqrt/div/log10/10^x per frame), "
                         "{ns} stereo streams x {sec:g} s @ 48 kHz per GPU, tilt automated every 512-frame block, in-place buffers"),
}


def peaks():
    p = REPO / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 100 ms while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.rows, self.proc, self.idx = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.idx)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = [float(r[1]) for r in self.rows if len(r) >= 8 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 8 and r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 8 for n, v in zip(names, r[4:8]) if v.lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def workload_params(W, ns, nf, first_stream=0):
    """Static [ns][n_sliders] parameters, or the per-block automation [n_blocks][ns][n_sliders] (host float32)."""
    import za_gpu
    from za_gpu import workloads as wl
    base = wl.default_params(W["plugin"], ns, first_stream)
    if not W["automate"]:
        return base
    sl = [dict(default=float(s.def_), min=float(s.min), max=float(s.max)) for s in za_gpu.slider_layout(W["plugin"])]
    nb = (nf + 511) // 512
    # trajectories of 64 seeded streams, tiled over the batch (the generator is a per-stream Python loop)
    k = min(ns, 64)
    a = wl.automation(sl, nb, k, which=W["automate"], first_stream=first_stream, base=base[0])
    return np.ascontiguousarray(np.tile(a, (1, (ns + k - 1) // k, 1))[:, :ns])


def host_input(W, n_streams, nf, first_stream=0):
    """Seeded host input of the workload's shape for the CPU legs: 8 generated streams tiled over the batch."""
    from za_gpu import workloads as wl
    k = min(n_streams, 8)
    if W["plugin"] == "cascade8":
        x = wl.noise_two_sines(k, nf, first_stream)
    else:
        x = wl.programme_bursts(k, nf, first_stream, n_ch=W["ch"][0])
    return np.ascontiguousarray(np.tile(x, ((n_streams + k - 1) // k, 1, 1))[:n_streams])


def cpu_reference_run(W, n_streams: int, seconds: float, threads: int):
    """Reference CPU path on a bounded sample of the workload -> (frames/s, kind, seconds taken)."""
    sys.path.insert(0, str(REPO / "oracle"))
    import za_oracle as zo
    orc = zo.best_oracle(W["plugin"])
    kind = "reference" if isinstance(orc, zo.RefPlugin) else "port"
    nf = int(seconds * SRATE)
    p = workload_params(W, n_streams, nf)
    x = host_input(W, n_streams, nf)
    orc.render(p[..., :threads, :] if p.ndim == 3 else p[:threads], x[:threads, :, :4800], threads=threads)          # warm
    t0 = time.perf_counter()
    orc.render(p, x, block=512, threads=threads)
    dt = time.perf_counter() - t0
    return n_streams * nf / dt, kind, dt


def workload_text(W, ns, sec):
    return W["text"].format(ns=ns, sec=sec)


def run_reference_arm(args, W):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    ns, secs = W["cpu"]
    vals = []
    for i in range(args.warmup + args.steps):
        v, kind, dt = cpu_reference_run(W, ns, secs, threads)
        if i >= args.warmup:
            vals.append((v, dt))
    v = float(np.mean([a for a, _ in vals]))
    sample = f"{ns} streams x {secs:g} s per step (of {W['streams']} x {W['seconds']:g} s), block 512, FTZ/DAZ on"
    line = {
        "impl": "reference", "metric": "aggregate sample-frames/s", "value": v, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": float(np.mean([d for _, d in vals]) * 1e3),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32" if W.get("f32_only") else "f64", "data": "synthetic",
        "x_realtime": v / SRATE,
        "config": {"workload": workload_text(W, W["streams"], W["seconds"]), "sample": sample},
        "cpu_baseline": {"value": v, "unit": "frames/s", "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def device_input(torch, W, ns, nf, seed):
    """Synthetic input generated on the device, [ns][ch][nf] float32.
    cascade8: uniform noise x0.25 + two sines per stream; dynamics / STFT: noise under a slow random
    envelope (-60..-6 dBFS in 50 ms knots) so gain computers traverse knee and hysteresis."""
    g = torch.Generator(device="cuda").manual_seed(seed)
    nch = W["ch"][0]
    x = torch.empty((ns, nch, nf), dtype=torch.float32, device="cuda")
    step = max(1, (256 << 20) // (nch * nf * 4))
    if W["plugin"] == "cascade8":
        t = torch.arange(nf, device="cuda", dtype=torch.float32) / SRATE
        f = torch.rand((ns, 2), device="cuda", generator=g)
        for s0 in range(0, ns, step):
            s1 = min(ns, s0 + step)
            blk = (torch.rand((s1 - s0, 1, nf), device="cuda", generator=g) - 0.5) * 0.5
            f1 = (60.0 + 340.0 * f[s0:s1, 0]).view(-1, 1, 1)
            f2 = (1000.0 + 5000.0 * f[s0:s1, 1]).view(-1, 1, 1)
            blk += 0.125 * torch.sin(2 * np.pi * f1 * t) + 0.125 * torch.sin(2 * np.pi * f2 * t)
            x[s0:s1] = blk
    else:
        seg = 2400
        nk = nf // seg + 2
        for s0 in range(0, ns, step):
            s1 = min(ns, s0 + step)
            knots = torch.rand((s1 - s0, 1, nk), device="cuda", generator=g) * 54.0 - 60.0
            env = torch.nn.functional.interpolate(knots, size=nk * seg, mode="linear", align_corners=False)[:, :, :nf]
            env = torch.pow(10.0, env / 20.0) * 0.7
            mono = torch.rand((s1 - s0, 1, nf), device="cuda", generator=g) * 2.0 - 1.0
            side = (torch.rand((s1 - s0, nch, nf), device="cuda", generator=g) * 2.0 - 1.0) * 0.3
            x[s0:s1] = env * (mono + side)
    torch.cuda.synchronize()
    return x


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="cfg2", choices=sorted(WORKLOADS) + ["cfg5"])
    ap.add_argument("--precision", default="f32", choices=["f32", "f64"])
    ap.add_argument("--streams", type=int, default=0)
    ap.add_argument("--seconds", type=float, default=0.0)
    ap.add_argument("--e2e-seconds", type=float, default=0.0, help="0 = the workload's length (cfg2) or 10 s (cfg3/cfg4)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-f64", action="store_true")
    ap.add_argument("--warps", type=int, default=0)
    args = ap.parse_args()
    if args.workload == "cfg5":
        import bench_cfg5
        bench_cfg5.main(args)
        return
    W = WORKLOADS[args.workload]
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    if args.impl == "reference":
        run_reference_arm(args, W)
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

    ns = args.streams or W["streams"]
    sec = args.seconds or W["seconds"]
    nf = int(sec * SRATE)
    n_in, n_out = W["ch"]
    bpf = 4 * (n_in + n_out)                      # algorithmic bytes per frame
    params = workload_params(W, ns, nf, first_stream=rank * ns)
    p0 = params if params.ndim == 2 else params[0]
    pauto = params if params.ndim == 3 else None
    x = device_input(torch, W, ns, nf, 5000 + rank)
    inplace = n_in == n_out and ns * (n_in + n_out) * nf * 4 > 120e9      # cfg3: 94 GB in + 94 GB out would not fit
    y = x if inplace else torch.empty((ns, n_out, nf), dtype=torch.float32, device="cuda")

    results = {}
    if W.get("f32_only") and args.precision != "f32":
        raise SystemExit(f"{args.workload}: float32 only (Faust leaf)")
    precs = [args.precision] + (["f64"] if args.precision == "f32" and not args.no_f64 and not W.get("f32_only") else [])
    for prec in precs:
        headline = prec == args.precision
        bp = za_gpu.BatchProcessor(W["plugin"], ns, SRATE, 512, prec, local)
        if args.warps:
            bp.set_option("warps_per_stream", args.warps)
        bp.set_stream(torch.cuda.current_stream().cuda_stream)
        bp.prepare(p0)
        steps = args.steps if headline else max(2, args.steps // 2)
        warm = args.warmup if headline else 1
        # parity spot check on the very first pass (fresh state): 4 streams x 1 s against the oracle
        pick = [0, ns // 3, ns // 2, ns - 1]
        xs = x[pick, :, :48000].cpu().numpy() if rank == 0 else None
        bp.process(x, out=y, params=pauto)
        torch.cuda.synchronize()
        parity = None
        if rank == 0:
            try:
                sys.path.insert(0, str(REPO / "oracle"))
                import za_oracle as zo
                pp = params[:94, pick] if params.ndim == 3 else params[pick]
                yr, yrd = zo.best_oracle(W["plugin"]).render(pp, xs, want_f64=True)
                got = y[pick, :, :48000].cpu().numpy()
                r = got.astype(np.float64) - yrd
                parity = {"streams": pick, "frames": 48000, "float_equal_frac": float(np.mean(got == yr)),
                          "residual_dbfs_rms": float(20 * np.log10(np.sqrt(np.mean(r ** 2)) + 1e-300))}
            except Exception as e:      # the checker is optional for the measurement
                parity = {"error": str(e)[:120]}
        for _ in range(warm - 1):
            bp.process(x, out=y, params=pauto)
        barrier()
        clk = ClockSampler(local)
        if headline and rank == 0:
            clk.start()
        l0 = bp.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            bp.process(x, out=y, params=pauto)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        kms, klaunch = bp.last_kernel_ms()     # dominant kernel of the LAST timed step (CUDA events on the stream)
        launches = bp.launch_count() - l0
        clocks = None
        if headline and rank == 0:
            # nvidia-smi samples every 100 ms: a timed region shorter than that is followed by identical, untimed
            # repeats of the same loop until ~0.8 s of load has been sampled (said so in the line)
            extra = 0
            t_ext = time.perf_counter()
            while len(clk.rows) < 6 and time.perf_counter() - t_ext < 1.5 and ms * steps < 600.0:
                for _ in range(steps):
                    bp.process(x, out=y, params=pauto)
                torch.cuda.synchronize()
                extra += steps
            clocks = clk.stop()
            clocks["window"] = "timed steps" if extra == 0 else f"timed steps + {extra} identical untimed steps (timed region shorter than the sampling period)"
        tmax = torch.tensor([ms], device="cuda")
        if dist is not None:
            dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        ms_max = float(tmax.item())
        results[prec] = dict(ms=ms_max, value=world * ns * nf / (ms_max * 1e-3), kernel_ms=kms, kernel_launches=klaunch,
                             launches=launches, clocks=clocks, parity=parity, steps=steps)
        bp.close()

    # ---- e2e: HOST buffers through the C ABI, copies inside the timed region ---------------------------
    e2e = None
    e_sec = args.e2e_seconds or (sec if args.workload == "cfg2" else min(sec, 10.0))
    enf = int(e_sec * SRATE)
    hx = hy = None
    try:
        hx = torch.empty((ns, n_in, enf), dtype=torch.float32, pin_memory=True)
        hy = torch.empty((ns, n_out, enf), dtype=torch.float32, pin_memory=True)
    except RuntimeError:
        hx = hy = None
    if hx is not None:
        hx.copy_(x[:, :, :enf])
        torch.cuda.synchronize()
        ep = None if pauto is None else np.ascontiguousarray(pauto[:(enf + 511) // 512])
        bp = za_gpu.BatchProcessor(W["plugin"], ns, SRATE, 512, args.precision, local).prepare(p0)
        for _ in range(2):
            bp.process(hx, out=hy, params=ep)
        barrier()
        t0 = time.perf_counter()
        e_steps = max(2, args.steps // 2)
        for _ in range(e_steps):
            bp.process(hx, out=hy, params=ep)   # returns when `hy` is complete
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / e_steps
        tm = torch.tensor([dt], device="cuda")
        if dist is not None:
            dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        dt = float(tm.item())
        e2e = {"value": world * ns * enf / dt, "unit": "frames/s", "h2d_bytes_per_step": ns * n_in * enf * 4,
               "d2h_bytes_per_step": ns * n_out * enf * 4, "ms_per_step": dt * 1e3, "streams_per_gpu": ns,
               "seconds_per_step": e_sec, "x_realtime": world * ns * enf / dt / SRATE, "host_memory": "pinned"}
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
    tp = REPO / "profiles" / "traffic.json"
    if tp.exists():
        try:
            traffic = json.loads(tp.read_text()).get(args.workload, {}).get(args.precision)
        except Exception:
            traffic = None

    def roof(rr, prec):
        if not rr["kernel_launches"] or rr["kernel_ms"] <= 0:
            return None
        alg = ns * nf * bpf                                     # algorithmic bytes of one step
        ach = alg / (rr["kernel_ms"] * 1e-3) / 1e9
        tr = None
        if traffic and traffic.get("dram_bytes_per_frame"):
            tr = traffic["dram_bytes_per_frame"] * ns * nf / rr["kernel_launches"]
        return {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "peak_source": peak_kind,
                "traffic": tr, "kernel": W["kernel"][prec], "kernel_ms_per_launch": rr["kernel_ms"] / rr["kernel_launches"],
                "launches_per_step": rr["kernel_launches"], "algorithmic_bytes_per_launch": alg // max(rr["kernel_launches"], 1),
                "algorithmic_bytes_per_frame": bpf}

    cpu = None
    if not args.no_cpu:
        threads = os.cpu_count() or 1
        cns, csec = W["cpu"]
        v, kind, dt = cpu_reference_run(W, cns, csec, threads)
        cpu = {"value": v, "unit": "frames/s", "cores": threads, "kind": kind,
               "sample": f"{cns} streams x {csec:g} s of the same workload ({dt:.1f} s of wall time), block 512",
               "x_realtime": v / SRATE}

    line = {
        "metric": "aggregate sample-frames/s", "value": r["value"], "unit": "frames/s", "n_gpus": world,
        "steps": r["steps"], "warmup": args.warmup, "ms_per_step": r["ms"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": args.precision, "data": "synthetic",
        "x_realtime": r["value"] / SRATE,
        "config": {"workload": workload_text(W, ns, sec),
                   "streams_per_gpu": ns, "frames_per_stream": nf, "block_len": 512, "io": "float32 planar [stream][ch][frame]",
                   "l2": "inputs (%.1f GB per step) larger than L2" % (ns * n_in * nf * 4 / 1e9), "parallelism": f"streams sharded x{world}"},
        "clocks": r["clocks"], "e2e": e2e, "gpu_launches": r["launches"], "roofline": roof(r, args.precision), "cpu_baseline": cpu,
        "parity_spot_check": r["parity"],
    }
    if "f64" in results and args.precision != "f64":
        q = results["f64"]
        line["f64"] = {"value": q["value"], "ms_per_step": q["ms"], "x_realtime": q["value"] / SRATE, "roofline": roof(q, "f64"),
                       "parity_spot_check": q["parity"], "gpu_launches": q["launches"]}
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def _only_json_on_stdout():
    """The contract is ONE JSON line on stdout.  Libraries print there too (NCCL's version banner under torchrun), so
    file descriptor 1 is pointed at stderr for the whole run and `print` keeps the real stdout."""
    sys.stdout.flush()
    real = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real, "w", buffering=1)


if __name__ == "__main__":
    _only_json_on_stdout()
    main()
