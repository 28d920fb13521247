#!/usr/bin/env python3
"""tools/make_profiles.py -- turn this round's gpurun_out/ captures into the tracked summaries under profiles/."""
import collections, csv, json, subprocess, sys
from pathlib import Path
REPO = Path(__file__).resolve().parent.parent
G, P = REPO / "gpurun_out", REPO / "profiles"
RND = "r01"
CAPS = [
    ("ee_v8", "easyexpander_split", "EasyExpander fp32 (`easyexpander_split_kernel`: band / filter / gain / helper warps)",
     "ncu --set full --clock-control none --import-source on -k regex:easyexpander_split -c 1 python tools/prof_run.py --plugin EasyExpander --streams 4096 --seconds 2 --reps 1",
     "4096 stereo streams x 96 000 frames; algorithmic bytes = 6.29 GB per launch"),
    ("stft_v4", "stft", "stft_mask2048 fp32 (`stft_mask_kernel<float,4>`)",
     "ncu --set full --clock-control none --import-source on -k regex:stft_mask -c 1 python tools/prof_run.py --plugin stft_mask2048 --streams 512 --seconds 4 --reps 1",
     "512 stereo streams x 192 000 frames; algorithmic bytes = 1.57 GB per launch"),
    ("gen_rtt_v6", "generic_rtt", "translated leaf RTT fp32 (`jsfx_process_kernel<g_RTT::Script,float>`)",
     "ncu --set full --clock-control none --import-source on -k regex:jsfx_process_kernel -c 1 python tools/prof_run.py --plugin RTT --streams 4096 --seconds 1 --reps 1",
     "4096 streams (8 in / 2 out) x 48 000 frames; algorithmic bytes = 7.86 GB per launch"),
    ("gen_ss_v4", "generic_spectralstabilizer", "translated leaf SpectralStabilizer fp32 (`jsfx_process_kernel<g_SpectralStabilizer::Script,float>`), a mem[]-heavy band loop",
     "ncu --set full --clock-control none --import-source on -k regex:jsfx_process_kernel -c 1 python tools/prof_run.py --plugin SpectralStabilizer --streams 4096 --seconds 0.5 --reps 1",
     "4096 stereo streams x 24 000 frames; algorithmic bytes = 1.57 GB per launch"),
    ("modtilt_v2", "modtilt", "Faust leaf ModTilt fp32 (`modtilt_kernel<8>`: one thread per stream, 8 streams per warp, stage-by-stage 16-frame runs)",
     "ncu --set full --clock-control none --import-source on -k regex:modtilt_kernel -c 1 python tools/prof_run.py --plugin ModTilt --streams 4096 --seconds 2 --reps 1",
     "4096 stereo streams x 96 000 frames; algorithmic bytes = 6.29 GB per launch"),
    ("c8_v3", "cascade8_x2_v3", "cascade8 fp32 (`cascade8_x2_kernel<2,1>`), re-captured at the end of the round",
     "ncu --set full --clock-control none --import-source on -k regex:cascade8_x2_kernel -c 1 python tools/prof_run.py --plugin cascade8 --streams 1024 --seconds 10 --reps 1",
     "1024 mono streams x 480 000 frames; algorithmic bytes = 3.93 GB per launch"),
]


def sh(*a):
    return subprocess.run(list(a), capture_output=True, text=True).stdout


def main():
    only = sys.argv[1:]            # e.g. `make_profiles.py modtilt`: that capture / launch list / bench line only, others untouched
    for rep, name, title, cmd, what in CAPS:
        if only and name not in only:
            continue
        f = G / f"{RND}_{rep}.ncu-rep"
        if not f.exists():
            print("missing", f); continue
        summ = sh(sys.executable, str(REPO / "tools/ncu_summary.py"), str(f))
        src = sh(sys.executable, str(REPO / "tools/ncu_source.py"), str(f), "12")
        raw = sh("ncu", "-i", str(f), "--page", "raw", "--csv")
        rows = list(csv.reader(raw.splitlines())); h, v = rows[0], rows[2]
        g = lambda k: v[h.index(k)] if k in h else "n/a"
        extra = "\n".join(f"    {k} = {g(k)}" for k in ("dram__bytes_read.sum", "dram__bytes_write.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
                                                       "smsp__inst_executed.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed"))
        body = f"# Round 1: {title}, ncu --set full\n\nCommand (after the same command exited 0 without ncu):\n`{cmd}`\n({what}.)\n\n## Summary\n\n" + \
               "\n".join("    " + l for l in summ.splitlines()) + "\n\n" + extra + "\n\n## Source page (warp-stall samples)\n\n" + "\n".join("    " + l for l in src.splitlines()) + "\n"
        if name == "easyexpander_split":
            roles = sh(sys.executable, str(REPO / "tools/ncu_roles.py"), str(f))
            body += "\n## Per warp role (tools/ncu_roles.py: SASS cut at every BAR.SYNC; a role's loop body = one segment, the short\nsegment after it = that role's wait at the batch barrier; order: band warps, filter warp, helper warps, gain warp)\n\n" + \
                    "\n".join("    " + l for l in roles.splitlines()) + "\n"
        (P / f"{RND}_{name}_ncu.md").write_text(body)
        print("wrote", P / f"{RND}_{name}_ncu.md")
    # launch lists: our kernels only (-k filter), share of the dominant kernel inside the benched command
    lines = []
    for w in (only or ("cfg2", "cfg3", "cfg4")):
        f = G / f"launches_{w}.csv"
        if not f.exists():
            continue
        txt = [l for l in f.read_text().splitlines() if not l.startswith("==")]
        (P / f"{RND}_launches_{w}.csv").write_text("\n".join(txt) + "\n")
        rows = list(csv.reader(txt)); h = rows[0]
        ki, vi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
        agg, n = collections.Counter(), collections.Counter()
        for r in rows[1:]:
            if len(r) <= vi:
                continue
            val = float(r[vi].replace(",", "")); u = r[ui]
            val = val / 1e6 if u.startswith("n") else (val / 1e3 if u.startswith("u") else val)
            k = r[ki].split("(")[0].replace("void ", "")
            agg[k] += val; n[k] += 1
        tot = sum(agg.values())
        lines.append(f"{w}: {sum(n.values())} launches of our kernels, {tot:.2f} ms under ncu (cold-cache, serialised)")
        for k, val in agg.most_common(6):
            lines.append(f"    {k:60s} n={n[k]:3d}  {val:10.3f} ms  share {100 * val / tot:5.1f} %")
    if only:
        with open(P / f"{RND}_launch_shares.txt", "a") as fh:
            fh.write("\n".join(lines) + "\n")
    else:
        (P / f"{RND}_launch_shares.txt").write_text("\n".join(lines) + "\n")
    print("\n".join(lines))
    # bench lines
    out = {}
    if only:
        for w in only:
            for tag in (w, "ref_" + w):
                f = G / f"bench_{tag}.json"
                if f.exists() and f.read_text().strip():
                    out[tag] = json.loads(f.read_text().strip().splitlines()[-1])
        (P / f"{RND}_bench_lines_{'_'.join(only)}.json").write_text(json.dumps(out, indent=1))
        print("bench lines:", list(out))
        return
    for w in ("ref_cfg2", "cfg2", "cfg3", "cfg4", "cfg5"):
        f = G / f"bench_{w}.json"
        if f.exists() and f.read_text().strip():
            out[w] = json.loads(f.read_text().strip().splitlines()[-1])
    (P / f"{RND}_bench_lines.json").write_text(json.dumps(out, indent=1))
    print("bench lines:", list(out))


if __name__ == "__main__":
    main()
