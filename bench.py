#!/usr/bin/env python3
"""bench.py -- BASELINE.json metric on the B200 batch renderer.

Headline at N=1 = BASELINE config 3, the "4096-stream workload" the north_star's >= 60 %-of-HBM target is stated on and
the largest configuration that fits one GPU: plugin `EasyExpander` (nonlinear envelope + gain computer), 4096 stereo
streams x 60 s at 48 kHz = 1.18e10 sample-frames per step, threshold / depth automated every 512-frame host block,
float32 planar I/O in place (94 GB; in + out would not fit 180 GB).  One step = one za_gpu_process() pass over the
whole batch; state carries from step to step as in a continuing render, and the programme material is REGENERATED on
the device between steps (outside the timed events), so every step renders the described input, not its own output.
The other BASELINE configs and leaf classes ride along as sub-records of the same JSON line (`workloads`): cfg2
(cascade8), cfg4 (stft_mask2048), cfg5 (mixed catalog), ModTilt / ERBTilt / ... (same shape as cfg3 on other leaves),
each with value, roofline, f64, e2e, cpu_baseline and parity_spot_check.  `--workload X` makes X the headline and
drops the sub-records; `--latency` prints the small-block latency report instead.

  value     frames/s with inputs resident in HBM (CUDA events on the launching stream around each step)
  f64       the same workload with fp64 arithmetic (what DSP-JSFX computes in; fp32 is the throughput mode)
  e2e       the same metric through the C ABI with HOST buffers: pinned input -> H2D -> kernels -> D2H -> pinned output
            inside the timed region (za_gpu_process(..., ZA_MEM_HOST)); `copy_ceiling` = the identical copy pattern
            with the kernels skipped
  roofline  algorithmic bytes (4 B x (in + out channels) per frame) / dominant-kernel time measured
            with CUDA events inside the timed region, against MEASURED_PEAKS.json hbm_gbs
  cpu_baseline  the reference's own compiled DSP-JSFX object (oracle/_ref; Faust leaves: the float restatement)
            multithreaded over a bounded sample of the same workload on this box's host cores
N>1: every rank renders its own batch on its own GPU (streams shard, no collective on the data path) -> weak scaling.
`--impl reference` times the CPU path alone (rank 0) on the same metric / config.  The repo arm's headline computes in
float32 (`dtype`), the CPU arm in float64 like the reference; the like-for-like fp64 GPU figure is `f64`.
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
SEED = 5000

# name -> plugin, streams per GPU, seconds, channels in/out, automated slider columns, CPU sample (streams, seconds), kernel,
# synth = input generator kind of za_gpu_synth (0 = noise + two sines, 1 = programme bursts)
WORKLOADS = {
    "cfg3": dict(plugin="EasyExpander", streams=4096, seconds=60.0, ch=(2, 2), automate=[0, 1], cpu=(256, 10.0), synth=1,
                 kernel={"f32": "easyexpander_v3_kernel", "f64": "easyexpander_kernel"},
                 text="cfg3: EasyExpander (nonlinear envelope + gain computer), {ns} stereo streams x {sec:g} s @ 48 kHz per GPU, "
                      "threshold/depth automated every 512-frame block, in-place buffers"),
    "cfg2": dict(plugin="cascade8", streams=1024, seconds=60.0, ch=(1, 1), automate=None, cpu=(1024, 20.0), synth=0,
                 kernel={"f32": "cascade8_x2_kernel", "f64": "cascade8_kernel"},
                 text="cfg2: cascade8 (8-section biquad/one-pole cascade), {ns} mono streams x {sec:g} s @ 48 kHz per GPU"),
    "cfg4": dict(plugin="stft_mask2048", streams=512, seconds=120.0, ch=(2, 2), automate=None, cpu=(64, 10.0), synth=1,
                 kernel={"f32": "stft_mask_kernel", "f64": "stft_mask_kernel"},
                 text="cfg4: stft_mask2048 (2048-pt Hann STFT, hop 512, per-bin gain, overlap-add), {ns} stereo streams x {sec:g} s @ 48 kHz per GPU"),
    # cfg3's shape on the Faust leaf of the Dynamics category (float32 only; CPU side = the restatement, parity unpinned)
    "modtilt": dict(plugin="ModTilt", streams=4096, seconds=60.0, ch=(2, 2), automate=[0], cpu=(512, 20.0), f32_only=True, synth=1,
                    kernel={"f32": "modtilt_kernel"},
                    text="cfg3 shape on the Faust leaf ModTilt (envelope tilt shaper: 5 one-poles + sqrt/div/log10/10^x per frame), "
                         "{ns} stereo streams x {sec:g} s @ 48 kHz per GPU, tilt automated every 512-frame block, in-place buffers"),
    # the other leaves of the three classes at the cfg3 batch shape (4096 stereo streams), 10 s per step
    "erbtilt": dict(plugin="ERBTilt", streams=4096, seconds=10.0, ch=(2, 2), automate=[0], cpu=(256, 5.0), synth=1,
                    kernel={"f32": "erbtilt_kernel", "f64": "erbtilt_kernel"},
                    text="ERB Tilt (15 one-poles x 2 ch + energy followers + guard), {ns} stereo streams x {sec:g} s @ 48 kHz per GPU, tilt automated every 512-frame block"),
    "specstab": dict(plugin="SpectralStabilizer", streams=4096, seconds=10.0, ch=(2, 2), automate=[1], cpu=(256, 5.0), synth=1,
                     kernel={"f32": "specstab_kernel", "f64": "specstab_kernel"},
                     text="Spectral Stabilizer (12 band-pass biquads x 2 ch + per-band gain smoothing), {ns} stereo streams x {sec:g} s @ 48 kHz per GPU, depth automated every 512-frame block"),
    "gts": dict(plugin="GTS", streams=4096, seconds=10.0, ch=(2, 2), automate=[0, 1], cpu=(64, 2.0), f32_only=True, synth=1,
                kernel={"f32": "gts_kernel"},
                text="Faust leaf GTS (257-tap Gaussian FIR rebuilt every sample from a smoothed sigma), {ns} stereo streams x {sec:g} s @ 48 kHz per GPU, sigma/attack automated every 512-frame block"),
    "var": dict(plugin="VAR", streams=4096, seconds=10.0, ch=(2, 2), automate=[0], cpu=(256, 5.0), f32_only=True, synth=1,
                kernel={"f32": "faust_leaf_kernel"},
                text="Faust leaf VAR (4 biquads x 2 ch + curvature detector + switching follower), {ns} stereo streams x {sec:g} s @ 48 kHz per GPU, amount automated every 512-frame block"),
    "rtt": dict(plugin="RTT", streams=4096, seconds=10.0, ch=(8, 2), automate=[0, 1], cpu=(256, 5.0), synth=1,
                kernel={"f32": "jsfx_process_kernel", "f64": "jsfx_process_kernel"},
                text="cfg3 on RTT (Reverb Tail Tamer, 8 in / 2 out), {ns} streams x {sec:g} s @ 48 kHz per GPU, amount/sensitivity automated every 512-frame block"),
    "soma": dict(plugin="SOMA", streams=4096, seconds=10.0, ch=(2, 2), automate=[0, 3], cpu=(256, 5.0), synth=1,
                 kernel={"f32": "jsfx_process_kernel", "f64": "jsfx_process_kernel"},
                 text="cfg3 on SOMA (lookahead limiter), {ns} stereo streams x {sec:g} s @ 48 kHz per GPU, drive/release automated every 512-frame block"),
}
SUBS_DEFAULT = ["cfg2", "cfg4", "modtilt", "erbtilt", "specstab", "cfg5"]


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


def host_input(W, streams, nf, t0=0):
    """The workload's input for the listed absolute stream numbers, on the CPU (bit-identical twin of the device generator)."""
    from za_gpu import workloads as wl
    return wl.synth(W["synth"], SEED, streams, W["ch"][0], nf, t0)


_CPU_INPUTS = {}


def cpu_reference_run(W, n_streams: int, seconds: float, threads: int):
    """Reference CPU path on a bounded sample of the workload -> (frames/s, kind, seconds taken)."""
    sys.path.insert(0, str(REPO / "oracle"))
    import za_oracle as zo
    orc = zo.best_oracle(W["plugin"])
    kind = "reference" if isinstance(orc, zo.RefPlugin) else "port"
    nf = int(seconds * SRATE)
    ck = (W["plugin"], n_streams, nf)
    if ck not in _CPU_INPUTS:                   # the same sample every step (inputs are not part of the timed region)
        _CPU_INPUTS.clear()
        p = workload_params(W, n_streams, nf)
        k = min(n_streams, 16)                  # 16 generated streams tiled over the sample (the generator twin is numpy)
        x = host_input(W, range(k), nf)
        _CPU_INPUTS[ck] = (p, np.ascontiguousarray(np.tile(x, ((n_streams + k - 1) // k, 1, 1))[:n_streams]))
    p, x = _CPU_INPUTS[ck]
    orc.render(p[..., :threads, :] if p.ndim == 3 else p[:threads], x[:threads, :, :4800], threads=threads)          # warm
    t0 = time.perf_counter()
    orc.render(p, x, block=512, threads=threads)
    dt = time.perf_counter() - t0
    return n_streams * nf / dt, kind, dt


def workload_text(W, ns, sec):
    return W["text"].format(ns=ns, sec=sec)


def config_of(W, ns, sec, world):
    nf = int(sec * SRATE)
    return {"workload": workload_text(W, ns, sec), "streams_per_gpu": ns, "frames_per_stream": nf, "block_len": 512,
            "io": "float32 planar [stream][ch][frame]",
            "l2": "inputs (%.1f GB per step) larger than L2" % (ns * W["ch"][0] * nf * 4 / 1e9),
            "parallelism": f"streams sharded x{world}"}


def run_reference_arm(args, W):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = int(os.environ.get("WORLD_SIZE", "1"))
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
        "dtype_note": "the reference computes DSP-JSFX in float64; the repo arm's headline is its float32 mode (its float64 figure is the line's `f64`)",
        "x_realtime": v / SRATE,
        "config": config_of(W, W["streams"], W["seconds"], world),
        "cpu_baseline": {"value": v, "unit": "frames/s", "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------------------------------
class Ctx:
    def __init__(self, torch, za_gpu, rank, world, local, dist):
        self.torch, self.za, self.rank, self.world, self.local, self.dist = torch, za_gpu, rank, world, local, dist
        self.peak, self.peak_kind = peaks()
        self.affinity0 = os.sched_getaffinity(0)
        self.numa = None

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, v: float) -> float:
        t = self.torch.tensor([v], device="cuda", dtype=self.torch.float64)
        if self.dist is not None:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())


def numa_bind(ctx):
    """Pin this rank's host threads (and so its pinned allocations) to the NUMA node of its GPU."""
    info = {"bound": False}
    try:
        pr = ctx.torch.cuda.get_device_properties(ctx.local)
        bus = f"{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
        node = int(Path(f"/sys/bus/pci/devices/{bus}/numa_node").read_text())
        info.update(pci=bus, node=node)
        if node >= 0:
            cpus = set()
            for part in Path(f"/sys/devices/system/node/node{node}/cpulist").read_text().strip().split(","):
                a, _, b = part.partition("-")
                cpus.update(range(int(a), int(b or a) + 1))
            target = cpus & ctx.affinity0
            info["allowed_cpus_on_node"] = len(target)
            if target:
                os.sched_setaffinity(0, target)
                info["bound"] = True
    except Exception as e:
        info["error"] = str(e)[:100]
    return info


def host_budget_bytes(world):
    try:
        for line in Path("/proc/meminfo").read_text().splitlines():
            if line.startswith("MemAvailable:"):
                return min(0.30 * int(line.split()[1]) * 1024 / world, 192e9)
    except Exception:
        pass
    return 24e9


def fill_host(ctx, W, buf, first_stream, enf):
    """Fill the pinned host array [ns][n_in][enf] with the workload's input, generated on the device in stream slabs."""
    torch = ctx.torch
    ns, n_in = buf.shape[0], buf.shape[1]
    slab = max(1, int((1 << 30) // (n_in * enf * 4)))
    for s0 in range(0, ns, slab):
        k = min(slab, ns - s0)
        t = torch.empty((k, n_in, enf), dtype=torch.float32, device="cuda")
        ctx.za.synth_into(t, W["synth"], SEED, first_stream + s0, cuda_stream=torch.cuda.current_stream().cuda_stream)
        torch.from_numpy(buf[s0:s0 + k]).copy_(t)
    torch.cuda.synchronize()


def measure(name, ctx, args, headline):
    """One workload on this rank's GPU -> record (rank 0 gets the full record, other ranks a stub)."""
    torch, za_gpu = ctx.torch, ctx.za
    W = WORKLOADS[name]
    rank, world, local = ctx.rank, ctx.world, ctx.local
    ns = (args.streams if headline and args.streams else W["streams"])
    sec = (args.seconds if headline and args.seconds else W["seconds"])
    nf = int(sec * SRATE)
    n_in, n_out = W["ch"]
    bpf = 4 * (n_in + n_out)                      # algorithmic bytes per frame
    first = rank * ns
    params = workload_params(W, ns, nf, first_stream=first)
    p0 = params if params.ndim == 2 else params[0]
    pauto = params if params.ndim == 3 else None
    stream_ptr = torch.cuda.current_stream().cuda_stream
    x = torch.empty((ns, n_in, nf), dtype=torch.float32, device="cuda")

    def regen():
        za_gpu.synth_into(x, W["synth"], SEED, first, cuda_stream=stream_ptr)

    regen()
    inplace = n_in == n_out and ns * (n_in + n_out) * nf * 4 > 120e9      # cfg3 shape: 94 GB in + 94 GB out would not fit
    y = x if inplace else torch.empty((ns, n_out, nf), dtype=torch.float32, device="cuda")

    results = {}
    prec0 = args.precision if headline else "f32"
    if W.get("f32_only") and prec0 != "f32":
        raise SystemExit(f"{name}: float32 only (Faust leaf)")
    precs = [prec0] + (["f64"] if prec0 == "f32" and not args.no_f64 and not W.get("f32_only") else [])
    for prec in precs:
        main_prec = prec == prec0
        bp = za_gpu.BatchProcessor(W["plugin"], ns, SRATE, 512, prec, local)
        if headline and args.warps:
            bp.set_option("warps_per_stream", args.warps)
        if headline:
            for ov in args.option:
                if ":" not in ov:                                   # "Leaf:name=value" belongs to bench_cfg5's leaves
                    bp.set_option(ov.split("=")[0], int(ov.split("=")[1]))
        bp.set_stream(stream_ptr)
        bp.prepare(p0)
        steps = (args.steps if headline else max(3, args.steps // 4)) if main_prec else max(2, args.steps // (2 if headline else 8))
        warm = (args.warmup if headline else 3) if main_prec else 1
        # parity spot check on the very first pass (fresh state): 4 streams x 94 whole host blocks (~1 s) against the oracle
        # (whole blocks: the oracle's render must not end in a short block the GPU render does not have -- samplesblock
        # is visible to scripts); the oracle's input is regenerated on the CPU by the generator's numpy twin
        pf = min(nf, 94 * 512)
        pick = sorted({0, ns // 3, ns // 2, ns - 1})
        bp.process(x, out=y, params=pauto)
        torch.cuda.synchronize()
        parity = None
        if rank == 0 and not args.no_parity:
            try:
                sys.path.insert(0, str(REPO / "oracle"))
                import za_oracle as zo
                xs = host_input(W, [first + s for s in pick], pf)
                pp = params[:94, pick] if params.ndim == 3 else params[pick]
                yr, yrd = zo.best_oracle(W["plugin"]).render(pp, xs, want_f64=True)
                got = y[pick, :, :pf].cpu().numpy()
                ref = yrd if yrd is not None else yr.astype(np.float64)
                r = got.astype(np.float64) - ref
                parity = {"streams": pick, "frames": pf, "float_equal_frac": float(np.mean(got == yr)),
                          "residual_dbfs_rms": float(20 * np.log10(np.sqrt(np.mean(r ** 2)) + 1e-300))}
            except Exception as e:      # the checker is optional for the measurement
                parity = {"error": str(e)[:160]}
        for _ in range(warm - 1):
            if inplace:
                regen()
            bp.process(x, out=y, params=pauto)
        if inplace:
            regen()
        ctx.barrier()
        clk = ClockSampler(local)
        sample_clocks = headline and main_prec and rank == 0
        if sample_clocks:
            clk.start()
        l0 = bp.launch_count()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        kms_sum, kl_sum = 0.0, 0
        for i in range(steps):
            ev[i][0].record()
            bp.process(x, out=y, params=pauto)
            ev[i][1].record()
            if inplace:
                regen()                            # outside the step's event pair: the next step renders fresh programme material
        torch.cuda.synchronize()
        ms = sum(a.elapsed_time(b) for a, b in ev) / steps
        kms, klaunch = bp.last_kernel_ms()     # dominant kernel of the LAST timed step (CUDA events on the stream)
        launches = bp.launch_count() - l0
        clocks = None
        if sample_clocks:
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
        ms_max = ctx.max_over_ranks(ms)
        results[prec] = dict(ms=ms_max, value=world * ns * nf / (ms_max * 1e-3), kernel_ms=kms, kernel_launches=klaunch,
                             launches=launches, clocks=clocks, parity=parity, steps=steps, warm=warm)
        bp.close()
    del x, y
    torch.cuda.empty_cache()

    # ---- e2e: HOST buffers through the C ABI, copies inside the timed region ---------------------------
    e2e = None
    if not args.no_e2e:
        budget = host_budget_bytes(world) if headline else min(host_budget_bytes(world), 26e9)
        e_sec = args.e2e_seconds if headline and args.e2e_seconds else sec
        per_sec = ns * (n_in + n_out) * 4 * SRATE
        if e_sec * per_sec > budget:
            e_sec = max(1.0, float(int(budget / per_sec)))
        enf = int(e_sec * SRATE) // 512 * 512
        hx = hy = None
        try:
            hx = za_gpu.HostBuffer((ns, n_in, enf))
            hy = za_gpu.HostBuffer((ns, n_out, enf))
            fill_host(ctx, W, hx.array, first, enf)
            ep = None if pauto is None else np.ascontiguousarray(pauto[:(enf + 511) // 512])
            bp = za_gpu.BatchProcessor(W["plugin"], ns, SRATE, 512, prec0, local).prepare(p0)
            e_warm, e_steps = (2, max(2, args.steps // 4)) if headline else (1, 2)
            for _ in range(e_warm):
                bp.process(hx.array, out=hy.array, params=ep)
            ctx.barrier()
            t0 = time.perf_counter()
            for _ in range(e_steps):
                bp.process(hx.array, out=hy.array, params=ep)   # returns when `hy` is complete
            dt = ctx.max_over_ranks((time.perf_counter() - t0) / e_steps)
            e2e = {"value": world * ns * enf / dt, "unit": "frames/s", "h2d_bytes_per_step": ns * n_in * enf * 4,
                   "d2h_bytes_per_step": ns * n_out * enf * 4, "ms_per_step": dt * 1e3, "steps": e_steps, "streams_per_gpu": ns,
                   "seconds_per_step": enf / SRATE, "x_realtime": world * ns * enf / dt / SRATE, "host_memory": "pinned (za_gpu_host_alloc)",
                   "numa": ctx.numa}
            if enf / SRATE < sec:
                e2e["note"] = (f"{enf / SRATE:g} s of the {sec:g} s per stream: the pinned in + out buffers are bounded to 30 % of "
                               f"the box's available host memory / {world} ranks (<= 192 GB per rank"
                               + ("" if headline else "; sub-records <= 26 GB") + ")")
            # the box's ceiling for this exact copy pattern: same buffers, same chunking, kernels skipped
            bp.set_option("copy_only", 1)
            bp.process(hx.array, out=hy.array, params=None)
            ctx.barrier()
            t0 = time.perf_counter()
            for _ in range(2):
                bp.process(hx.array, out=hy.array, params=None)
            dtc = ctx.max_over_ranks((time.perf_counter() - t0) / 2)
            e2e["copy_ceiling"] = {"value": world * ns * enf / dtc, "unit": "frames/s", "ms_per_step": dtc * 1e3,
                                   "h2d_gbs_per_gpu": ns * n_in * enf * 4 / dtc / 1e9, "d2h_gbs_per_gpu": ns * n_out * enf * 4 / dtc / 1e9,
                                   "frac_of_ceiling": dtc / dt}
            bp.close()
        except Exception as e:
            e2e = {"error": str(e)[:200]}
        finally:
            for h in (hx, hy):
                if h is not None:
                    h.free()

    if rank != 0:
        return None

    r = results[prec0]
    traffic = None
    tp = REPO / "profiles" / "traffic.json"
    if tp.exists():
        try:
            traffic = json.loads(tp.read_text()).get(name, {})
        except Exception:
            traffic = None

    def roof(rr, prec):
        if not rr["kernel_launches"] or rr["kernel_ms"] <= 0:
            return None
        alg = ns * nf * bpf                                     # algorithmic bytes of one step
        ach = alg / (rr["kernel_ms"] * 1e-3) / 1e9
        tr = None
        t = (traffic or {}).get(prec)
        if t and t.get("dram_bytes_per_frame"):
            tr = t["dram_bytes_per_frame"] * ns * nf / rr["kernel_launches"]
        return {"bound": "hbm", "achieved": ach, "peak": ctx.peak, "unit": "GB/s", "frac": ach / ctx.peak, "peak_source": ctx.peak_kind,
                "traffic": tr, "kernel": W["kernel"].get(prec), "kernel_ms_per_launch": rr["kernel_ms"] / rr["kernel_launches"],
                "launches_per_step": rr["kernel_launches"], "algorithmic_bytes_per_launch": alg // max(rr["kernel_launches"], 1),
                "algorithmic_bytes_per_frame": bpf}

    cpu = None
    if not args.no_cpu:
        os.sched_setaffinity(0, ctx.affinity0)
        threads = os.cpu_count() or 1
        cns, csec = W["cpu"]
        if not headline:
            csec = max(2.0, csec / 4)
        try:
            v, kind, dt = cpu_reference_run(W, cns, csec, threads)
            cpu = {"value": v, "unit": "frames/s", "cores": threads, "kind": kind,
                   "sample": f"{cns} streams x {csec:g} s of the same workload ({dt:.1f} s of wall time), block 512",
                   "x_realtime": v / SRATE}
        except Exception as e:
            cpu = {"error": str(e)[:160]}

    rec = {
        "value": r["value"], "unit": "frames/s", "ms_per_step": r["ms"], "steps": r["steps"], "warmup": r["warm"], "dtype": prec0,
        "x_realtime": r["value"] / SRATE, "config": config_of(W, ns, sec, world),
        "clocks": r["clocks"], "e2e": e2e, "gpu_launches": r["launches"], "roofline": roof(r, prec0), "cpu_baseline": cpu,
        "parity_spot_check": r["parity"],
        "input": "regenerated on the device before every step (in-place buffers)" if inplace else "device-resident, out of place",
    }
    if "f64" in results and prec0 != "f64":
        q = results["f64"]
        rec["f64"] = {"value": q["value"], "ms_per_step": q["ms"], "steps": q["steps"], "x_realtime": q["value"] / SRATE, "roofline": roof(q, "f64"),
                      "parity_spot_check": q["parity"], "gpu_launches": q["launches"]}
    return rec


def latency_report(ctx, calls, plugins=("EasyExpander",)):
    """Small-block realtime latency (north_star: measured and reported, not optimised): ONE za_gpu_process call of
    64 / 128 / 256 / 512 frames through pinned HOST buffers (H2D + kernels + D2H + synchronise), wall clock per call."""
    za_gpu = ctx.za
    out = {"unit": "us", "calls": calls, "what": "one za_gpu_process(..., ZA_MEM_HOST) call = one host block; p50 / p99 over the calls"}
    for key in plugins:
        info = za_gpu.plugin_info(key)
        rows = {}
        for ns in (1, 64, 4096):
            for nfr in (64, 128, 256, 512):
                bp = za_gpu.BatchProcessor(key, ns, SRATE, nfr, "f32", ctx.local).prepare(None)
                hx, hy = za_gpu.HostBuffer((ns, info.n_in, nfr)), za_gpu.HostBuffer((ns, info.n_out, nfr))
                hx.array[:] = (np.random.default_rng(1).random(hx.shape, np.float32) - 0.5) * 0.2
                for _ in range(10):
                    bp.process(hx.array, out=hy.array)
                t = np.empty(calls)
                for i in range(calls):
                    t0 = time.perf_counter()
                    bp.process(hx.array, out=hy.array)
                    t[i] = time.perf_counter() - t0
                rows[f"{ns}x{nfr}"] = {"p50": float(np.percentile(t, 50) * 1e6), "p99": float(np.percentile(t, 99) * 1e6),
                                       "block_us_at_48k": nfr / SRATE * 1e6}
                bp.close(); hx.free(); hy.free()
        out[key] = rows
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=None, choices=sorted(WORKLOADS) + ["cfg5"])
    ap.add_argument("--subs", default=None, help="comma list of sub-record workloads ('' = none); default: all when --workload is not given")
    ap.add_argument("--precision", default="f32", choices=["f32", "f64"])
    ap.add_argument("--streams", type=int, default=0)
    ap.add_argument("--seconds", type=float, default=0.0)
    ap.add_argument("--e2e-seconds", type=float, default=0.0, help="0 = the workload's length, bounded by the host-memory budget")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-f64", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--latency", action="store_true", help="print the small-block latency report (1000 calls per shape) and exit")
    ap.add_argument("--warps", type=int, default=0)
    ap.add_argument("--option", action="append", default=[], help="name=value passed to za_gpu_set_option on the headline handle (experiments)")
    args = ap.parse_args()
    head = args.workload or "cfg3"
    subs = ([] if args.workload else SUBS_DEFAULT) if args.subs is None else [s for s in args.subs.split(",") if s]
    if head == "cfg5":
        import bench_cfg5
        bench_cfg5.main(args)
        return
    W = WORKLOADS[head]
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
    ctx = Ctx(torch, za_gpu, rank, world, local, dist)
    ctx.numa = numa_bind(ctx)

    if args.latency:
        if rank == 0:
            print(json.dumps({"metric": "small-block latency", "latency": latency_report(ctx, 1000, ("EasyExpander", "cascade8"))}), flush=True)
        if dist is not None:
            dist.barrier(); dist.destroy_process_group()
        return

    t_start = time.perf_counter()
    rec = measure(head, ctx, args, headline=True)
    sub_recs = {}
    for name in subs:
        t1 = time.perf_counter()
        try:
            if name == "cfg5":
                import bench_cfg5
                r = bench_cfg5.run(args, ctx, sub=True)
            else:
                r = measure(name, ctx, args, headline=False)
        except Exception as e:                      # a sub-record must never take the headline down
            r = {"error": f"{type(e).__name__}: {str(e)[:200]}"}
            torch.cuda.empty_cache()
        if rank == 0 and r is not None:
            r["wall_s"] = time.perf_counter() - t1
            sub_recs[name] = r
    lat = None
    if rank == 0 and not args.workload and not args.no_e2e:
        try:
            os.sched_setaffinity(0, ctx.affinity0)
            lat = latency_report(ctx, 200)
        except Exception as e:
            lat = {"error": str(e)[:160]}

    if rank == 0:
        line = {
            "metric": "aggregate sample-frames/s", "value": rec["value"], "unit": "frames/s", "n_gpus": world,
            "steps": rec["steps"], "warmup": rec["warmup"], "ms_per_step": rec["ms_per_step"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": rec["dtype"], "data": "synthetic",
            "dtype_note": "headline = the renderer's float32 mode (the throughput mode the HBM target is stated on); `f64` = the same step in float64, "
                          "the arithmetic the reference's DSP-JSFX path (and the --impl reference arm) computes in",
            "x_realtime": rec["x_realtime"], "config": rec["config"], "clocks": rec["clocks"], "e2e": rec["e2e"],
            "gpu_launches": rec["gpu_launches"], "roofline": rec["roofline"], "cpu_baseline": rec["cpu_baseline"],
            "parity_spot_check": rec["parity_spot_check"], "input": rec["input"],
        }
        if "f64" in rec:
            line["f64"] = rec["f64"]
        if sub_recs:
            line["workloads"] = sub_recs
        if lat is not None:
            line["latency"] = lat
        line["bench_wall_s"] = time.perf_counter() - t_start
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
