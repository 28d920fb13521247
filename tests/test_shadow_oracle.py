"""The SECONDARY oracle (BASELINE.json configs[0]): the reference's own WDL EEL2 interpreter running the
plugin source (oracle/shadow_eel.cpp -> oracle/_ref/libshadow_eel.so, a JUCE-free restatement of the reference's
--correctness-check shadow VM) against the reference's compiled DSP-JSFX path (oracle/_ref/libref_<key>.so behind
oracle/ref_shim.c).  It pins the HARNESS: that ref_shim.c pushes sliders, mirrors aliases and orders @init / @slider /
@block / @sample the way the host does.  Tolerances as the reference ships them (JSFXCorrectnessCheck.h:34-35:
1e-5 audio, 1e-8 scalar); the measured agreement is far tighter and is asserted at SURVEY 8c's table values x4.

Needs the plugin sources (/root/reference), so it runs in the build container, not on the GPU box."""
import os
from pathlib import Path

import numpy as np
import pytest

import za_oracle as zo
from za_gpu import workloads as wl

REF = Path(os.environ.get("ZA_REFERENCE_ROOT", "/root/reference"))
AUDIO_EPS, SCALAR_EPS = 1.0e-5, 1.0e-8            # kAudioCompareEpsilon / kScalarCompareEpsilon


def _pair(key):
    import build_ref
    src_path = build_ref.PLUGINS[key][0]
    if not (zo.REF_DIR / "libshadow_eel.so").exists() or not zo.have_ref(key) or not src_path.exists():
        pytest.skip("needs oracle/_ref (compiled reference + shadow EEL2) and the plugin source")
    ref = zo.RefPlugin(key)
    sh = zo.ShadowPlugin(key, src_path.read_text(errors="replace"), ref.n_in, ref.n_out, sliders=ref.sliders)
    return ref, sh


def test_cfg1_ddt_compiled_vs_shadow():
    """BASELINE config 1 as worded: DDT, 10 s stereo 48 kHz, the compiled DSP-JSFX path with --correctness-check =
    compiled vs shadow EEL2, block 512, declared defaults, the survey's gated-noise + sine input."""
    ref, sh = _pair("DDT")
    x = wl.cfg1_input()
    p = np.array([[s["default"] for s in ref.sliders]], np.float32)
    assert list(p[0]) == [30, 50, 40, 55, 2, 100, 0, 0, 50]
    y, yd, v = ref.render(p, x, want_f64=True, want_vars=True)
    ys, yds = sh.render_one(p[0], x[0])
    d = np.abs(yds - yd[0])
    assert d.max() <= AUDIO_EPS
    assert d.max() == 0.0                            # bit-identical over 960 000 values
    assert np.array_equal(ys, y[0])
    checked = 0
    for name, val in v.items():
        if name.startswith("__fnlocal__"):
            continue                                 # per-(section, function) locals exist only in the compiled path
        sv = sh.var(name)
        if np.isnan(sv) and np.isnan(val[0]):
            continue
        assert abs(sv - val[0]) <= SCALAR_EPS * max(1.0, abs(val[0])), name
        checked += 1
    assert checked > 20
    sh.close()


# key, seconds, max |delta| allowed (SURVEY 8c's measured agreement of the reference's two paths, x4 head-room)
TABLE = [("RTT", 5, 0.0), ("ATTACK", 5, 0.0), ("SOMA", 5, 0.0), ("EasyExpander", 5, 4.4e-16), ("ERBTilt", 2, 1.8e-15),
         ("SpectralStabilizer", 2, 6.4e-15), ("BedRock", 2, 1.4e-13), ("cascade8", 2, 1e-12), ("stft_mask2048", 1, 1e-14)]


@pytest.mark.parametrize("key,sec,tol", TABLE)
def test_compiled_vs_shadow_table(key, sec, tol):
    ref, sh = _pair(key)
    nf = int(sec * 48000)
    x = wl.noise_two_sines(1, nf) if key == "cascade8" else wl.programme_bursts(1, nf, n_ch=ref.n_in)
    p = wl.cascade8_params(1) if key == "cascade8" else np.array([[s["default"] for s in ref.sliders]], np.float32)
    y, yd = ref.render(p, x, want_f64=True)
    ys, yds = sh.render_one(p[0], x[0])
    d = np.abs(yds - yd[0]).max()
    print(f"compiled vs shadow EEL2 {key}: max |delta| {d:.3e}, float outputs differing {int(np.sum(ys != y[0]))} of {ys.size}")
    assert d <= AUDIO_EPS
    assert d <= tol, d
    sh.close()


def test_shadow_follows_block_automation():
    """Slider pushes per host block: the shadow's @slider runs when a quantised value changed, like the compiled path's."""
    ref, sh = _pair("EasyExpander")
    nf = 512 * 40 + 100
    x = wl.programme_bursts(1, nf)
    p = wl.automation(ref.sliders, (nf + 511) // 512, 1, which=[0, 1, 2])
    y, yd = ref.render(p, x, want_f64=True)
    ys, yds = sh.render_one(p[:, 0], x[0])
    assert np.abs(yds - yd[0]).max() <= 4.4e-16
    sh.close()
