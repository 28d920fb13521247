#!/usr/bin/env python3
"""tools/make_profiles_r02.py -- round 2: turn gpurun_out/r02_*.ncu-rep (ncu --set full, one launch each) into the tracked
summaries profiles/r02_<kernel>_ncu.md and refresh profiles/traffic.json with the measured DRAM bytes per stream-frame."""
import csv, json, subprocess, sys
from pathlib import Path
REPO = Path(__file__).resolve().parent.parent
G, P = REPO / "gpurun_out", REPO / "profiles"
# report, summary name, title, command, frames in the captured launch, traffic.json key
CAPS = [
    ("r02_modtilt_scan", "modtilt_scan", "Faust leaf ModTilt fp32, time-parallel scan kernel (`modtilt_scan_kernel`: one warp per stream, 256-frame tiles; correctly rounded maps, before the MUFU forms)",
     "python bench.py --workload modtilt --seconds 4 --steps 1 --warmup 3 --no-cpu --no-e2e --no-parity", 4096 * 24000, None),
    ("r02_erbtilt", "erbtilt", "ERB Tilt fp32, hand-written scan kernel (`erbtilt_kernel<float, 8>`)",
     "python bench.py --workload erbtilt --seconds 2 --steps 1 --warmup 3 --no-cpu --no-e2e --no-parity --no-f64", 4096 * 96000, "erbtilt"),
    ("r02_specstab", "specstab", "Spectral Stabilizer fp32, hand-written scan kernel (`specstab_kernel<float, 8>`)",
     "python bench.py --workload specstab --seconds 2 --steps 1 --warmup 3 --no-cpu --no-e2e --no-parity --no-f64", 4096 * 96000, "specstab"),
    ("r02_modtilt_fast", "modtilt_scan_fast", "Faust leaf ModTilt fp32, scan kernel with the MUFU forms of sqrt / div / log10 / 10^x (the default)",
     "python bench.py --workload modtilt --seconds 4 --steps 1 --warmup 3 --no-cpu --no-e2e --no-parity", None, "modtilt"),
    ("r02_easyexpander", "easyexpander_v3", "EasyExpander fp32, v3 role pipeline (`easyexpander_v3_kernel`: 4 band warps, filter, G1, G2 and 14 active helper warps per CTA)",
     "python bench.py --workload cfg3 --seconds 2 --steps 1 --warmup 3 --no-cpu --no-e2e --no-parity --no-f64 --subs ''", 4096 * 96000, "cfg3"),
    ("r02_cascade8", "cascade8_x2", "cascade8 fp32 (`cascade8_x2_kernel`)",
     "python bench.py --workload cfg2 --seconds 10 --steps 1 --warmup 3 --no-cpu --no-e2e --no-parity --no-f64", 1024 * 480000, "cfg2"),
    ("r02_cbg_tiled", "cbg_tiled", "Faust leaf ClickBeGoneSG fp32, tiled kernel (`cbg_tiled_kernel`: 8 FIR warps with lane = frame + 1 recurrence warp with lane = stream per 32-stream CTA, inputs by cp.async one tile ahead)",
     "python tools/prof_run.py --plugin ClickBeGoneSG --streams 4096 --seconds 2 --reps 2", 4096 * 96000, None),
    ("r02_soma_staged", "soma_staged", "translated leaf SOMA fp32 as a 6-stage pipeline (`jsfx_staged_kernel`: one warp per group of @sample statements, shared-memory rings between them, rand() state through a cp.async window)",
     "python tools/prof_run.py --plugin SOMA --streams 5041 --seconds 1 --reps 2", 5041 * 48000, None),
    ("r02_rtt_staged", "rtt_staged", "translated leaf RTT fp32 (8 inputs, 2 outputs) as an 8-stage pipeline (`jsfx_staged_kernel`, 2-frame batches)",
     "python tools/prof_run.py --plugin RTT --streams 5041 --seconds 1 --reps 2", 5041 * 48000, None),
    ("r02_red_tiled", "red_tiled", "Faust leaf RED fp32 (6 in / 6 out), tiled kernel (`red_tiled_kernel`: 8 map warps lane = frame + 2 recurrence warps lane = stream per 32-stream CTA)",
     "python tools/prof_run.py --plugin RED --streams 5041 --seconds 1 --reps 2", 5041 * 48000, None),
    ("r02_var_tiled", "var_tiled", "Faust leaf VAR fp32, tiled kernel (`var_tiled_kernel`: 4 map warps lane = frame + 3 recurrence warps lane = stream per 16-stream CTA)",
     "python tools/prof_run.py --plugin VAR --streams 5041 --seconds 1 --reps 2", 5041 * 48000, None),
    ("r02_gts", "gts", "Faust leaf GTS fp32 (`gts_kernel`: time-parallel 257-tap Gaussian FIR rebuilt per frame)",
     "python bench.py --workload gts --seconds 2 --steps 1 --warmup 3 --no-cpu --no-e2e --no-parity", 4096 * 96000, "gts"),
]


def sh(*a):
    return subprocess.run(list(a), capture_output=True, text=True).stdout


def num(s):
    return float(s.replace(",", ""))


def main():
    tj = P / "traffic.json"
    traffic = json.loads(tj.read_text()) if tj.exists() else {}
    for rep, name, title, cmd, frames, tkey in CAPS:
        f = G / f"{rep}.ncu-rep"
        if not f.exists():
            continue
        summ = sh(sys.executable, str(REPO / "tools/ncu_summary.py"), str(f))
        src = sh(sys.executable, str(REPO / "tools/ncu_source.py"), str(f), "14")
        raw = sh("ncu", "-i", str(f), "--page", "raw", "--csv")
        rows = list(csv.reader(raw.splitlines())); h, u, v = rows[0], rows[1], rows[2]
        g = lambda k: v[h.index(k)] if k in h else "n/a"
        unit = lambda k: u[h.index(k)] if k in h else ""
        extra = "\n".join(f"    {k} [{unit(k)}] = {g(k)}" for k in ("dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum",
                                                                   "sm__throughput.avg.pct_of_peak_sustained_elapsed"))
        note = ""
        if frames and g("dram__bytes_read.sum") != "n/a":
            scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}
            b = num(g("dram__bytes_read.sum")) * scale.get(unit("dram__bytes_read.sum"), 1.0) + num(g("dram__bytes_write.sum")) * scale.get(unit("dram__bytes_write.sum"), 1.0)
            inst = num(g("smsp__inst_executed.sum")) if g("smsp__inst_executed.sum") != "n/a" else 0
            note = f"\n    => {b / frames:.2f} DRAM bytes per stream-frame (algorithmic 16, cfg2: 8); {inst / frames:.2f} warp-instructions = {32 * inst / frames:.0f} thread-slots per frame\n"
            if tkey:
                traffic.setdefault(tkey, {})["f32"] = {"dram_bytes_per_frame": round(b / frames, 3), "source": f"profiles/r02_{name}_ncu.md: {b / 1e9:.3f} GB over {frames} stream-frames"}
        body = (f"# Round 2: {title}, ncu --set full --clock-control none\n\nCommand (after the same command exited 0 without ncu): `{cmd}`\n\n## Summary\n\n"
                + "\n".join("    " + l for l in summ.splitlines()) + "\n\n" + extra + "\n" + note + "\n## Source page (warp-stall samples)\n\n"
                + "\n".join("    " + l for l in src.splitlines()) + "\n")
        (P / f"r02_{name}_ncu.md").write_text(body)
        print("wrote", P / f"r02_{name}_ncu.md", note.strip())
    tj.write_text(json.dumps(traffic, indent=1))


if __name__ == "__main__":
    main()
