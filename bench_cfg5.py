"""bench_cfg5.py -- BASELINE config 5 (`bench.py --workload cfg5`): mixed-catalog batch bounce.

65,536 streams drawn round-robin from the leaves this library renders (hand-written classes: cascade8,
EasyExpander, stft_mask2048; translated leaves: RTT, SOMA, ATTACK, ERBTilt, SpectralStabilizer -- the five Faust
leaves of SURVEY.md section 8d are not built and are left out, said so in `config`), sharded over the GPUs of the
run (streams / N per GPU, strong scaling over a fixed batch is NOT what is measured: every rank renders
65,536 / N streams, so the whole-job batch is fixed and `scaling` is "strong").  One step = one 1 s time chunk
of every stream (the 10 min bounce is 600 such chunks with state carried; inputs are generated on the device
per leaf, outside the timed region, and reused).  Every leaf has its own handle and CUDA stream, so the eight
kernels of a step overlap on the device inside the two phases described at LEAVES below; the step is timed with
CUDA events bracketing all of them.
"""
from __future__ import annotations

import json
import os
import sys
import time
from pathlib import Path

import numpy as np

REPO = Path(__file__).resolve().parent
SRATE = 48000.0
# Schedule of one step, two phases, every leaf on its own CUDA stream inside a phase:
#   A: the hand-written leaves and the heap-less translated leaves (RTT, ATTACK) -- short kernels with big shared-memory
#      tiles;
#   B: the translated leaves that keep their state in a mem[] heap (SpectralStabilizer, ERBTilt, SOMA) -- long kernels
#      that live off the L1.  Shared memory is carved out of the same 228 KB as the L1, so while phase-A kernels are
#      resident on an SM the heap leaves' lines are evicted (measured: SpectralStabilizer / ERBTilt 3x slower inside a
#      free-for-all mix, 600-650 ms per step; catalog order 578, longest first 510, + stream priorities 479-646).
LEAVES = ["stft_mask2048", "EasyExpander", "cascade8", "RTT", "ATTACK", "SpectralStabilizer", "ERBTilt", "SOMA"]
PHASE_B = {"SpectralStabilizer", "ERBTilt", "SOMA"}
NOT_BUILT = ["VAR", "ModTilt", "ClickBeGoneSG", "GTS", "RED"]
TOTAL_STREAMS = 65536


def main(args):
    import torch
    import bench as B
    import za_gpu
    from za_gpu import workloads as wl

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        if rank == 0:
            reference_arm(args)
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the renderer has no CPU fallback")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    total = args.streams or TOTAL_STREAMS
    per_leaf = total // world // len(LEAVES)
    sec = args.seconds or 1.0
    nf = int(sec * SRATE)
    prec = args.precision
    jobs, alg_bytes, frames = [], 0, 0
    for li, key in enumerate(LEAVES):
        info = za_gpu.plugin_info(key)
        W = dict(plugin=key, ch=(info.n_in, info.n_out))
        st = torch.cuda.Stream()
        bp = za_gpu.BatchProcessor(key, per_leaf, SRATE, 512, prec, local)
        bp.set_stream(st.cuda_stream)
        p = wl.default_params(key, per_leaf, first_stream=rank * per_leaf)
        bp.prepare(p)
        x = B.device_input(torch, W, per_leaf, nf, 7000 + 13 * rank + li)
        y = x if info.n_in == info.n_out else torch.empty((per_leaf, info.n_out, nf), dtype=torch.float32, device="cuda")
        jobs.append((key, bp, st, x, y, info))
        alg_bytes += per_leaf * nf * 4 * (info.n_in + info.n_out)
        frames += per_leaf * nf
    torch.cuda.synchronize()

    sa = [j[2] for j in jobs if j[0] not in PHASE_B]
    sb = [j[2] for j in jobs if j[0] in PHASE_B]

    def step():
        for a_ in sa:                                # phase A starts when the previous step's phase B is done
            for b_ in sb:
                a_.wait_stream(b_)
        for key, bp, _, x, y, _ in jobs:
            if key not in PHASE_B:
                bp.process(x, out=y)
        for b_ in sb:                                # phase B
            for a_ in sa:
                b_.wait_stream(a_)
        for key, bp, _, x, y, _ in jobs:
            if key in PHASE_B:
                bp.process(x, out=y)

    def fence(ev_stream):
        for _, _, st, _, _, _ in jobs:          # the timing stream waits for every leaf's stream
            ev_stream.wait_stream(st)

    main_stream = torch.cuda.current_stream()
    warm = max(args.warmup, 3)
    for _ in range(warm):
        step()
    torch.cuda.synchronize()
    barrier()
    clk = B.ClockSampler(local)
    if rank == 0:
        clk.start()
    l0 = sum(j[1].launch_count() for j in jobs)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _, _, st, _, _, _ in jobs:
        st.wait_stream(main_stream)
    e0.record(main_stream)
    for _, _, st, _, _, _ in jobs:
        st.wait_event(e0)
    for _ in range(args.steps):
        step()
    fence(main_stream)
    e1.record(main_stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    launches = sum(j[1].launch_count() for j in jobs) - l0
    per_leaf_ms = {k: bp.last_kernel_ms()[0] for k, bp, *_ in jobs}
    clocks = clk.stop() if rank == 0 else None
    tmax = torch.tensor([ms], device="cuda")
    if dist is not None:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    ms = float(tmax.item())

    # ---- e2e: host buffers through the C ABI for every leaf (a 1/8 slice of the streams), copies timed -----------
    e_n = max(32, per_leaf // 8)
    e_jobs = []
    for key, _, _, x, _, info in jobs:
        hx = torch.empty((e_n, info.n_in, nf), dtype=torch.float32, pin_memory=True)
        hx.copy_(x[:e_n])
        hy = torch.empty((e_n, info.n_out, nf), dtype=torch.float32, pin_memory=True)
        bp = za_gpu.BatchProcessor(key, e_n, SRATE, 512, prec, local).prepare(wl.default_params(key, e_n))
        e_jobs.append((bp, hx, hy))
    torch.cuda.synchronize()
    for bp, hx, hy in e_jobs:
        bp.process(hx, out=hy)
    barrier()
    t0 = time.perf_counter()
    for bp, hx, hy in e_jobs:
        bp.process(hx, out=hy)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    tm = torch.tensor([dt], device="cuda")
    if dist is not None:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
    dt = float(tm.item())
    e_frames = world * e_n * nf * len(LEAVES)
    e2e = {"value": e_frames / dt, "unit": "frames/s", "ms_per_step": dt * 1e3,
           "h2d_bytes_per_step": int(sum(h.numel() * 4 for _, h, _ in e_jobs)), "d2h_bytes_per_step": int(sum(h.numel() * 4 for _, _, h in e_jobs)),
           "streams_per_gpu": e_n * len(LEAVES), "seconds_per_step": sec, "host_memory": "pinned",
           "note": "leaves rendered one after another through za_gpu_process(..., ZA_MEM_HOST)"}
    for bp, _, _ in e_jobs:
        bp.close()
    if rank != 0:
        if dist is not None:
            dist.barrier(); dist.destroy_process_group()
        return
    peak, peak_kind = B.peaks()
    value = world * frames / (ms * 1e-3)
    ach = world * alg_bytes / (ms * 1e-3) / 1e9 / world       # per GPU
    cpu = None if args.no_cpu else cpu_sample()
    line = {
        "metric": "aggregate sample-frames/s", "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
        "warmup": warm, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": prec,
        "data": "synthetic", "x_realtime": value / SRATE,
        "config": {"workload": f"cfg5: mixed-catalog bounce, {total} streams round-robin over {len(LEAVES)} leaves "
                               f"({', '.join(LEAVES)}), one step = a {sec:g} s time chunk of every stream (the 10 min bounce = 600 chunks, state carried)",
                   "leaves_not_built": NOT_BUILT, "streams_per_gpu": per_leaf * len(LEAVES), "streams_per_leaf_per_gpu": per_leaf,
                   "frames_per_stream_per_step": nf, "block_len": 512, "l2": "inputs (%.1f GB per step per GPU) larger than L2" % (alg_bytes / 2e9),
                   "parallelism": f"streams sharded x{world}; per step two phases (hand-written + heap-less leaves, then the mem[]-heap leaves), one CUDA stream per leaf inside a phase"},
        "clocks": clocks, "e2e": e2e, "gpu_launches": launches,
        "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "peak_source": peak_kind,
                     "traffic": None, "kernel": "all leaves' kernels of one step, overlapped", "kernel_ms_per_launch": ms,
                     "algorithmic_bytes_per_launch": alg_bytes, "per_leaf_kernel_ms": per_leaf_ms},
        "cpu_baseline": cpu,
    }
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier(); dist.destroy_process_group()


def cpu_sample(ns=32, sec=1.0):
    """The reference's compiled DSP for every leaf of the mix, multithreaded, on a bounded sample."""
    sys.path.insert(0, str(REPO / "oracle"))
    sys.path.insert(0, str(REPO / "zorakaudio-experimental-plugins_b200"))
    import za_oracle as zo
    from za_gpu import workloads as wl
    threads = os.cpu_count() or 1
    nf = int(sec * SRATE)
    frames, t = 0, 0.0
    for key in LEAVES:
        orc = zo.best_oracle(key)
        x = wl.noise_two_sines(ns, nf) if key == "cascade8" else wl.programme_bursts(ns, nf, n_ch=orc.n_in)
        p = wl.cascade8_params(ns) if key == "cascade8" else np.tile(np.array([s["default"] for s in orc.sliders], np.float32), (ns, 1))
        orc.render(p[:threads], x[:threads, :, :4800], threads=threads)
        t0 = time.perf_counter()
        orc.render(p, x, block=512, threads=threads)
        t += time.perf_counter() - t0
        frames += ns * nf
    return {"value": frames / t, "unit": "frames/s", "cores": threads, "kind": "reference",
            "sample": f"{ns} streams x {sec:g} s of each of the {len(LEAVES)} leaves ({t:.1f} s of wall time), block 512"}


def reference_arm(args):
    vals = []
    for i in range(args.warmup + args.steps):
        c = cpu_sample()
        if i >= args.warmup:
            vals.append(c)
    v = float(np.mean([c["value"] for c in vals]))
    c = dict(vals[-1], value=v)
    print(json.dumps({"impl": "reference", "metric": "aggregate sample-frames/s", "value": v, "unit": "frames/s", "n_gpus": args.gpus,
                      "steps": args.steps, "warmup": args.warmup, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                      "dtype": "f64", "data": "synthetic", "config": {"workload": "cfg5 mixed-catalog bounce", "sample": c["sample"]},
                      "cpu_baseline": c, "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}), flush=True)
