#!/usr/bin/env python3
"""Generate tests/golden/*.npz by running the REFERENCE's own compiled DSP (oracle/_ref, built from
/root/reference by oracle/build_ref.py) on seeded inputs.  Run in the build container only."""
import hashlib
import sys
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
REPO = HERE.parent.parent
sys.path.insert(0, str(REPO / "oracle"))
sys.path.insert(0, str(REPO / "zorakaudio-experimental-plugins_b200"))
import za_oracle as zo  # noqa: E402
from za_gpu import workloads as wl  # noqa: E402

EXCERPT = 2048


def golden(key, params, x, block=512, tag=""):
    ref = zo.RefPlugin(key)
    y, yd = ref.render(params, x, block=block, want_f64=True, threads=1)
    out = HERE / f"{key}{tag}.npz"
    np.savez_compressed(
        out, params=np.asarray(params, np.float32), block=block,
        n_streams=x.shape[0], n_frames=x.shape[2],
        y64_head=yd[:, :, :EXCERPT], y64_tail=yd[:, :, -EXCERPT:],
        y32_sha256=np.frombuffer(hashlib.sha256(y.tobytes()).digest(), np.uint8),
        y64_sha256=np.frombuffer(hashlib.sha256(yd.tobytes()).digest(), np.uint8),
        x_sha256=np.frombuffer(hashlib.sha256(x.tobytes()).digest(), np.uint8),
        y64_abs_sum=np.abs(yd).sum(axis=2))
    print(out, out.stat().st_size)


if __name__ == "__main__":
    ns, nf = 8, 48000
    golden("cascade8", wl.cascade8_params(ns), wl.noise_two_sines(ns, nf))
    # automation: parameters move every 512-frame block
    base = wl.cascade8_params(4)
    nb = 20
    auto = np.repeat(base[None], nb, axis=0).copy()
    auto[:, :, 4] *= np.linspace(1.0, 2.0, nb)[:, None]      # sweep the S3 band frequency
    auto[5:, :, 14] = 0.25                                   # and step the mix at block 5
    golden("cascade8", auto, wl.noise_two_sines(4, nb * 512 - 100), tag="_auto")
    # EasyExpander (cfg3 shape): threshold / depth automated every block, programme-like bursts
    ee = zo.RefPlugin("EasyExpander")
    nf = 48000
    st = zo.RefPlugin("stft_mask2048")
    nfs = 24000 + 77
    golden("stft_mask2048", wl.automation(st.sliders, (nfs + 511) // 512, 3, which=[0, 1, 4]), wl.programme_bursts(3, nfs), tag="_auto")
    golden("EasyExpander", wl.automation(ee.sliders, (nf + 511) // 512, 6, which=[0, 1]), wl.programme_bursts(6, nf), tag="_auto")
    # translated leaves (csrc/gen): one stream, 9000 frames, first three sliders automated every block
    leaves = {}
    for key in ("RTT", "SOMA", "ATTACK", "ERBTilt", "SpectralStabilizer", "DDT", "DPT", "SaliencePush", "ADS", "BedRock",
                "Roomalizer", "TSEQ"):
        ref = zo.RefPlugin(key)
        nfl = 9000
        base = np.array([s["default"] for s in ref.sliders], np.float32)
        p = wl.automation(ref.sliders, (nfl + 511) // 512, 1, which=list(range(min(3, len(base)))), base=base)
        x = wl.programme_bursts(1, nfl, n_ch=ref.n_in)
        y, yd = ref.render(p, x, want_f64=True, threads=1)
        leaves[key + "_y64_tail"] = yd[0, :2, -1024:]
        leaves[key + "_y64_sha256"] = np.frombuffer(hashlib.sha256(yd.tobytes()).digest(), np.uint8)
    np.savez_compressed(HERE / "translated_leaves.npz", **leaves)
    print(HERE / "translated_leaves.npz", (HERE / "translated_leaves.npz").stat().st_size)
