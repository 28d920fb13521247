"""bench_cfg5.py -- BASELINE config 5 (`bench.py --workload cfg5`, and the `cfg5` sub-record of the default line):
mixed-catalog batch bounce.

65,536 streams drawn round-robin from the 13 leaves SURVEY.md section 8(d) names for cfg5 (those this library renders;
any leaf still missing is listed in `config.leaves_not_built`), sharded over the GPUs of the run: every rank renders
65,536 / N streams, so the whole-job batch is fixed and `scaling` is "strong".  One step = one 1 s time chunk of every
stream (the 10 min bounce is 600 such chunks with state carried).  As SURVEY 8(d) prescribes, inputs are generated ON
THE DEVICE from the counter-based generator (za_gpu_synth, seed = absolute stream number, absolute time = chunk index)
outside the timed region, outputs are reduced to per-stream checksums + energy by a device kernel, and a random
subset of streams of every leaf is re-rendered by the CPU oracle (oracle/_ref, or the float restatement for Faust
leaves) for the first chunk: `parity_spot_check`.  Every leaf has its own handle and CUDA stream, so the kernels of a
step overlap on the device as described at PHASE_B below; the step is timed with CUDA events
bracketing all of them.
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
SEED = 7000
# The cfg5 leaf set of SURVEY.md section 8(d), in its order.
CFG5_LEAVES = ["EasyExpander", "RTT", "SOMA", "ATTACK", "ERBTilt", "SpectralStabilizer", "cascade8", "stft_mask2048",
               "VAR", "ModTilt", "ClickBeGoneSG", "GTS", "RED"]
# Schedule of one step (`--option sched=N`; every leaf renders on its own CUDA stream):
#   5 (default)  no phases: all 13 leaves at once, the long latency-bound stream-per-thread kernels launched first, SOMA / RTT /
#                VAR on high-priority CUDA streams.  Measured on B200 after ERBTilt / SpectralStabilizer / ClickBeGoneSG got
#                their own kernels: 165 ms per step against 257 ms for the two-phase schedule on the same box (SOMA takes
#                105 ms alone and ~150 ms inside the mix, and it is what the step now waits for).  With the stage pipelines
#                (SOMA / RTT / ATTACK), the tiled RED / VAR kernels and the next chunk's generators kept out of the timed
#                region: 118 ms, and the schedule hardly matters any more.
#   4            as 5 without stream priorities (187-190 ms then, 119 now); `--option order=K` picks another launch order
#                (179-213 ms then, 118-126 now).
#   0            the round-1 two-phase schedule: A = everything but PHASE_B, then B = the translated leaf that keeps its
#                state in a mem[] heap (SOMA) alone -- shared memory is carved out of the same 228 KB as the L1, and while the
#                heap leaves SpectralStabilizer / ERBTilt were still translated they ran 3x slower inside a free-for-all mix.
#   1            0 with ClickBeGoneSG outside the phase barrier (596-626 ms against 505-520 ms when it was still the generic
#                kernel with its delay lines in local memory).
#   2 / 3        0 with VAR + RED, or with the scan kernels, moved into phase B (within 2 % of 0).
PHASE_B = {"SOMA"}
LONG_POLE = {"ClickBeGoneSG"}
PHASE_B_WIDE = {"SOMA", "VAR", "RED"}
PHASE_B_SCAN = {"SOMA", "ERBTilt", "SpectralStabilizer", "ModTilt"}
LAUNCH_FIRST = ["ClickBeGoneSG", "RTT", "RED", "ATTACK", "VAR"]
LAUNCH_ORDERS = [LAUNCH_FIRST,
                 ["GTS", "SOMA", "RTT", "ATTACK", "VAR", "ClickBeGoneSG", "RED", "SpectralStabilizer", "ERBTilt", "stft_mask2048", "EasyExpander", "ModTilt", "cascade8"],   # longest solo first
                 ["stft_mask2048", "EasyExpander", "GTS", "SpectralStabilizer", "ERBTilt", "SOMA", "RTT", "ATTACK", "VAR", "ClickBeGoneSG", "RED"],   # whole-SM kernels first
                 ["SOMA", "RTT", "ATTACK", "GTS", "VAR", "RED", "ClickBeGoneSG", "stft_mask2048", "EasyExpander", "SpectralStabilizer", "ERBTilt"],
                 ["SOMA", "RTT", "VAR", "RED", "ATTACK", "ClickBeGoneSG", "GTS"],
                 ["EasyExpander", "cascade8", "stft_mask2048", "ModTilt", "ERBTilt", "SpectralStabilizer", "GTS", "ClickBeGoneSG"],   # bandwidth kernels first
                 ["RTT", "SOMA", "ATTACK", "RED", "VAR"],
                 ["SOMA", "ClickBeGoneSG", "RTT", "RED", "ATTACK", "VAR"]]
TOTAL_STREAMS = 65536
FAUST = {"VAR", "ModTilt", "ClickBeGoneSG", "GTS", "RED"}


def leaf_oracle(key):
    sys.path.insert(0, str(REPO / "oracle"))
    import za_oracle as zo
    if key in FAUST:
        return zo.PortPlugin(key), "port (float32 restatement, parity unpinned)"
    o = zo.best_oracle(key)
    return o, ("reference" if isinstance(o, zo.RefPlugin) else "port")


def run(args, ctx, sub=False):
    torch, za_gpu = ctx.torch, ctx.za
    from za_gpu import workloads as wl
    rank, world, local = ctx.rank, ctx.world, ctx.local
    have = set(za_gpu.plugin_keys())
    leaves = [k for k in CFG5_LEAVES if k in have]
    for ov in getattr(args, "option", None) or []:          # --option only=A+B+C: a subset of the leaves (experiments; not cfg5)
        if ov.startswith("only="):
            leaves = [k for k in leaves if k in ov[5:].split("+")]
    not_built = [k for k in CFG5_LEAVES if k not in have]
    total = (args.streams if not sub and args.streams else TOTAL_STREAMS)
    per_leaf = total // world // len(leaves)
    sec = (args.seconds if not sub and args.seconds else 1.0)
    nf = int(sec * SRATE)
    prec = "f32" if sub else args.precision
    steps = max(3, args.steps // 4) if sub else args.steps
    hi_prio = not any(ov.startswith("sched=") and ov != "sched=5" for ov in (getattr(args, "option", None) or []))
    jobs, alg_bytes, frames = [], 0, 0
    for li, key in enumerate(leaves):
        info = za_gpu.plugin_info(key)
        st = torch.cuda.Stream(priority=-1) if (hi_prio and key in ("SOMA", "RTT", "VAR")) else torch.cuda.Stream()
        p_leaf = "f32" if info.kind == 1 else prec
        bp = za_gpu.BatchProcessor(key, per_leaf, SRATE, 512, p_leaf, local)
        bp.set_stream(st.cuda_stream)
        for ov in getattr(args, "option", None) or []:          # --option Leaf:name=value (experiments on one leaf of the mix)
            if ov.startswith(key + ":") or ov.startswith("*:"):
                name, val = ov.split(":", 1)[1].split("=")
                bp.set_option(name, int(val))
        first = (rank * len(leaves) + li) * per_leaf            # absolute stream numbers of this leaf's shard
        p = wl.default_params(key, per_leaf, first_stream=first)
        bp.prepare(p)
        x = torch.empty((per_leaf, info.n_in, nf), dtype=torch.float32, device="cuda")
        kind = 0 if key == "cascade8" else 1
        y = x if info.n_in == info.n_out else torch.empty((per_leaf, info.n_out, nf), dtype=torch.float32, device="cuda")
        jobs.append(dict(key=key, bp=bp, st=st, x=x, y=y, info=info, first=first, kind=kind, p=p))
        alg_bytes += per_leaf * nf * 4 * (info.n_in + info.n_out)
        frames += per_leaf * nf
    torch.cuda.synchronize()

    def gen(chunk):                                # inputs of time chunk `chunk`, on every leaf's own stream
        for j in jobs:
            za_gpu.synth_into(j["x"], j["kind"], SEED, j["first"], t0=chunk * nf, cuda_stream=j["st"].cuda_stream)

    sched = 5
    for ov in getattr(args, "option", None) or []:
        if ov.startswith("sched="):
            sched = int(ov.split("=")[1])
    # launch order inside phase A: the long latency-bound kernels first (they hold one or two warps per SM for the whole
    # phase), the short bandwidth-heavy hand-written kernels after them
    first = LAUNCH_FIRST
    for ov in getattr(args, "option", None) or []:          # --option order=K: launch-order experiments
        if ov.startswith("order="):
            first = LAUNCH_ORDERS[int(ov.split("=")[1])]
    order = sorted(jobs, key=lambda j: first.index(j["key"]) if j["key"] in first else len(first)) if sched else jobs
    free = LONG_POLE if sched == 1 else set()
    phase_b = PHASE_B_WIDE if sched == 2 else (PHASE_B_SCAN if sched == 3 else (set() if sched >= 4 else PHASE_B))   # 4, 5: no phases
    sa = [j["st"] for j in jobs if j["key"] not in phase_b and j["key"] not in free]
    sb = [j["st"] for j in jobs if j["key"] in phase_b]

    def step():
        for a_ in sa:                                # phase A starts when the previous step's phase B is done
            for b_ in sb:
                a_.wait_stream(b_)
        for j in order:
            if j["key"] not in phase_b:
                j["bp"].process(j["x"], out=j["y"])
        for b_ in sb:                                # phase B: after phase A, except the long pole, which it overlaps
            for a_ in sa:
                b_.wait_stream(a_)
        for j in order:
            if j["key"] in phase_b:
                j["bp"].process(j["x"], out=j["y"])

    main_stream = torch.cuda.current_stream()

    def fence():
        for j in jobs:          # the timing stream waits for every leaf's stream
            main_stream.wait_stream(j["st"])

    # ---- chunk 0 from fresh state: checksums + sampled parity against the CPU oracle ---------------------------------
    gen(0)
    step()
    torch.cuda.synchronize()
    parity, sums = {}, {}
    if rank == 0 and not args.no_parity:
        rng = np.random.default_rng(12345)
        n_pick = 16 if not sub else 4
        for j in jobs:
            key = j["key"]
            try:
                cs, en = za_gpu.checksum(j["y"])
                sums[key] = {"checksum_xor": int(np.bitwise_xor.reduce(cs.ravel())), "energy": float(en.sum())}
                pick = sorted(rng.choice(per_leaf, size=min(n_pick, per_leaf), replace=False).tolist())
                orc, okind = leaf_oracle(key)
                xs = wl.synth(j["kind"], SEED, [j["first"] + s for s in pick], j["info"].n_in, nf, 0)
                yr, yrd = orc.render(j["p"][pick], xs, want_f64=True)
                got = j["y"][pick].cpu().numpy()
                ref = yrd if yrd is not None else yr.astype(np.float64)
                r = got.astype(np.float64) - ref
                # the CPU twin of the checksum on the oracle's float output: equal iff every sampled output bit is
                ccs, _ = wl.checksum(yr)
                parity[key] = {"streams": len(pick), "frames": nf, "oracle": okind,
                               "residual_dbfs_rms": float(20 * np.log10(np.sqrt(np.mean(r ** 2)) + 1e-300)),
                               "float_equal_frac": float(np.mean(got == yr)),
                               "checksum_equal_frac": float(np.mean(ccs == cs[pick]))}
            except Exception as e:
                parity[key] = {"error": f"{type(e).__name__}: {str(e)[:120]}"}

    warm = 3 if sub else max(args.warmup, 3)
    for i in range(warm - 1):
        gen(1 + i)
        step()
    torch.cuda.synchronize()
    ctx.barrier()
    import bench as B
    clk = B.ClockSampler(local)
    if rank == 0 and not sub:
        clk.start()
    l0 = sum(j["bp"].launch_count() for j in jobs)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    host_issue_ms = 0.0
    for i in range(steps):
        if i:                                       # the generators of chunk i must not run under step i - 1's last kernels: each
            for j in jobs:                          # leaf's stream would start them as soon as ITS leaf is done, inside the
                j["st"].wait_event(ev[i - 1][1])    # previous step's events (measured: 145 ms per step with that overlap, 118-125 without)
        gen(warm + i)                               # the next time chunk of every stream (outside the step's events)
        fence()
        ev[i][0].record(main_stream)
        for j in jobs:
            j["st"].wait_event(ev[i][0])
        t_h = time.perf_counter()
        step()
        host_issue_ms += (time.perf_counter() - t_h) * 1e3 / steps      # host time to issue the step (the GPU is not waited for)
        if i == steps - 1:
            for j in jobs:                          # when each leaf's stream ran dry in the last step (timeline of the mix)
                j["ev_end"] = torch.cuda.Event(enable_timing=True)
                j["ev_end"].record(j["st"])
        fence()
        ev[i][1].record(main_stream)
    torch.cuda.synchronize()
    ms = sum(a.elapsed_time(b) for a, b in ev) / steps
    launches = sum(j["bp"].launch_count() for j in jobs) - l0
    per_leaf_ms = {j["key"]: j["bp"].last_kernel_ms()[0] for j in jobs}
    # [start, end] of each leaf's kernels inside the last step, ms after the step's start event (end = the stream's end event,
    # start = end - the kernels' own duration)
    per_leaf_span = {j["key"]: [round(ev[-1][0].elapsed_time(j["ev_end"]) - per_leaf_ms[j["key"]], 1), round(ev[-1][0].elapsed_time(j["ev_end"]), 1)] for j in jobs}
    clocks = clk.stop() if rank == 0 and not sub else None
    ms = ctx.max_over_ranks(ms)

    # ---- e2e: host buffers through the C ABI for every leaf (a 1/8 slice of the streams), copies timed -----------
    e2e = None
    if not args.no_e2e:
        e_n = max(32, per_leaf // 8)
        e_jobs = []
        for j in jobs:
            info = j["info"]
            hx, hy = za_gpu.HostBuffer((e_n, info.n_in, nf)), za_gpu.HostBuffer((e_n, info.n_out, nf))
            torch.from_numpy(hx.array).copy_(j["x"][:e_n])
            bp = za_gpu.BatchProcessor(j["key"], e_n, SRATE, 512, "f32" if info.kind == 1 else prec, local).prepare(j["p"][:e_n])
            e_jobs.append((bp, hx, hy))
        torch.cuda.synchronize()
        for bp, hx, hy in e_jobs:
            bp.process(hx.array, out=hy.array)
        ctx.barrier()
        t0 = time.perf_counter()
        for bp, hx, hy in e_jobs:
            bp.process(hx.array, out=hy.array)
        torch.cuda.synchronize()
        dt = ctx.max_over_ranks(time.perf_counter() - t0)
        e_frames = world * e_n * nf * len(leaves)
        e2e = {"value": e_frames / dt, "unit": "frames/s", "ms_per_step": dt * 1e3,
               "h2d_bytes_per_step": int(sum(h.array.size * 4 for _, h, _ in e_jobs)), "d2h_bytes_per_step": int(sum(h.array.size * 4 for _, _, h in e_jobs)),
               "streams_per_gpu": e_n * len(leaves), "seconds_per_step": sec, "host_memory": "pinned (za_gpu_host_alloc)",
               "note": "leaves rendered one after another through za_gpu_process(..., ZA_MEM_HOST)"}
        for bp, hx, hy in e_jobs:
            bp.close(); hx.free(); hy.free()
    for j in jobs:
        j["bp"].close()
    jobs.clear()
    torch.cuda.empty_cache()
    if rank != 0:
        return None
    value = world * frames / (ms * 1e-3)
    ach = alg_bytes / (ms * 1e-3) / 1e9                        # per GPU
    cpu = None
    if not args.no_cpu:
        os.sched_setaffinity(0, ctx.affinity0)
        cpu = cpu_sample(leaves, *( (64, 1.0) if sub else (1024, 30.0) ))
    return {
        "metric": "aggregate sample-frames/s", "value": value, "unit": "frames/s", "n_gpus": world, "steps": steps,
        "warmup": warm, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": prec,
        "data": "synthetic", "x_realtime": value / SRATE,
        "config": {"workload": f"cfg5: mixed-catalog bounce, {total} streams round-robin over {len(leaves)} leaves "
                               f"({', '.join(leaves)}), one step = a {sec:g} s time chunk of every stream (the 10 min bounce = 600 chunks, state carried)",
                   "leaves_not_built": not_built, "streams_per_gpu": per_leaf * len(leaves), "streams_per_leaf_per_gpu": per_leaf,
                   "frames_per_stream_per_step": nf, "block_len": 512, "l2": "inputs (%.1f GB per step per GPU) larger than L2" % (alg_bytes / 2e9),
                   "input": "generated on the device per time chunk by za_gpu_synth (counter-based, seed = absolute stream number), outside the timed events",
                   "parallelism": f"streams sharded x{world}; one CUDA stream per leaf, schedule {sched} (bench_cfg5.py: 5 = all leaves at once, long latency-bound kernels launched first and on high-priority streams; 0 = two phases)"},
        "clocks": clocks, "e2e": e2e, "gpu_launches": launches,
        "roofline": {"bound": "hbm", "achieved": ach, "peak": ctx.peak, "unit": "GB/s", "frac": ach / ctx.peak, "peak_source": ctx.peak_kind,
                     "traffic": None, "kernel": "all leaves' kernels of one step, overlapped", "kernel_ms_per_launch": ms,
                     "algorithmic_bytes_per_launch": alg_bytes, "per_leaf_kernel_ms": per_leaf_ms, "host_issue_ms_per_step": host_issue_ms, "step_ms": [round(a.elapsed_time(b), 1) for a, b in ev], "per_leaf_span_ms_last_step": per_leaf_span},
        "cpu_baseline": cpu, "parity_spot_check": parity, "output_checksums_chunk0": sums,
    }


def main(args):
    import torch
    import bench as B
    import za_gpu
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        if rank == 0:
            reference_arm(args)
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the renderer has no CPU fallback")
    torch.cuda.set_device(local)
    torch.cuda.set_stream(torch.cuda.Stream())
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = B.Ctx(torch, za_gpu, rank, world, local, dist)
    line = run(args, ctx, sub=False)
    if rank == 0:
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier(); dist.destroy_process_group()


def built_leaves():
    sys.path.insert(0, str(REPO / "zorakaudio-experimental-plugins_b200"))
    import za_gpu
    have = set(za_gpu.plugin_keys())
    return [k for k in CFG5_LEAVES if k in have]


def cpu_sample(leaves, ns=1024, sec=30.0):
    """The reference's compiled DSP (Faust leaves: the restatement) for every leaf of the mix, multithreaded, on a bounded
    sample: `ns` streams x `sec` seconds split evenly over the leaves."""
    sys.path.insert(0, str(REPO / "zorakaudio-experimental-plugins_b200"))
    from za_gpu import workloads as wl
    threads = os.cpu_count() or 1
    nf = int(sec * SRATE)
    per = max(threads, ns // len(leaves))
    frames, t = 0, 0.0
    for key in leaves:
        orc, _ = leaf_oracle(key)
        k = min(per, 16)
        x = wl.synth(0 if key == "cascade8" else 1, SEED, range(k), orc.n_in, nf)
        x = np.ascontiguousarray(np.tile(x, ((per + k - 1) // k, 1, 1))[:per])
        p = wl.default_params(key, per)
        orc.render(p[:threads], x[:threads, :, :4800], threads=threads)
        t0 = time.perf_counter()
        orc.render(p, x, block=512, threads=threads)
        t += time.perf_counter() - t0
        frames += per * nf
    return {"value": frames / t, "unit": "frames/s", "cores": threads, "kind": "reference (Faust leaves: restatement)",
            "sample": f"{per} streams x {sec:g} s of each of the {len(leaves)} leaves ({t:.1f} s of wall time), block 512"}


def reference_arm(args):
    leaves = built_leaves()
    vals = []
    for i in range(args.warmup + args.steps):
        c = cpu_sample(leaves, 256, 4.0)
        if i >= args.warmup:
            vals.append(c)
    v = float(np.mean([c["value"] for c in vals]))
    c = dict(vals[-1], value=v)
    print(json.dumps({"impl": "reference", "metric": "aggregate sample-frames/s", "value": v, "unit": "frames/s", "n_gpus": args.gpus,
                      "steps": args.steps, "warmup": args.warmup, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                      "dtype": "f64", "data": "synthetic", "config": {"workload": "cfg5 mixed-catalog bounce", "sample": c["sample"]},
                      "cpu_baseline": c, "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}), flush=True)
