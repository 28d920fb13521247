"""CPU tests of the float restatements of the Faust leaves RED, ClickBeGoneSG, VAR and GTS (oracle/za_faust_port.c; PARITY
UNPINNED: no Faust compiler here, no Faust render in the reference's tests).  They pin the restatements to properties the
.dsp files imply, so the GPU parity tests (tests/test_zfaust_leaves_gpu.py) compare against something that is at least
structurally right."""
import numpy as np
import pytest

import za_oracle as zo
from za_gpu import workloads as wl


def defaults(port, ns):
    return np.tile(np.array([s["default"] for s in port.sliders], np.float32), (ns, 1))


def test_gts_is_a_128_frame_delay_when_attack_equals_sustain():
    """shaped = aGain (x@128 - sustain) + sGain sustain (GTS.dsp:84-88): equal gains -> x@128 times that gain, whatever sigma."""
    port = zo.PortPlugin("GTS")
    ns, nf = 2, 24000
    x = wl.programme_bursts(ns, nf)
    p = defaults(port, ns)
    p[:, 0] = [0.5, 6.0]
    y, _ = port.render(p, x)
    # the five si.smoo start at 0 (mix, out gain ramp in; time constant 1088 frames, and a float smoother stalls within
    # eps / (1 - p) = 6.5e-5 of its target): after they have settled the output is the delayed input
    assert np.abs(y[:, :, 20000:] - x[:, :, 20000 - 128:nf - 128]).max() < 1e-4 * np.abs(x).max()
    # sustain raised by 6 dB on a DC input: the Gaussian has unit DC gain (:31) -> output = DC * 10^(6/20)
    p[:, 2] = 6.0
    dc = np.full((ns, 2, nf), 0.25, np.float32)
    y, _ = port.render(p, dc)
    assert np.abs(y[:, :, -1] - 0.25 * 10 ** (6 / 20)).max() < 2e-4


def test_clickbegone_passes_clean_material_with_15_frames_of_latency_and_repairs_a_click():
    port = zo.PortPlugin("ClickBeGoneSG")
    nf = 9000
    t = np.arange(nf) / 48000.0
    x = np.zeros((1, 2, nf), np.float32)
    x[0, 0] = 0.2 * np.sin(2 * np.pi * 220 * t)
    x[0, 1] = 0.2 * np.sin(2 * np.pi * 330 * t)
    p = defaults(port, 1)
    y, _ = port.render(p, x)
    # xC = L@15 (:83-84), mix = 0 without a trigger (the onset of the sines itself trips the detector once: base starts at 0)
    assert np.abs(y[0, :, 215:] - x[0, :, 200:-15]).max() < 1e-6
    xc = x.copy()
    xc[0, :, 5000] += 0.8                                                 # a needle click
    yc, _ = port.render(p, xc)
    assert abs(yc[0, 0, 5015] - xc[0, 0, 5000]) > 0.3                     # replaced towards the SG prediction
    assert abs(yc[0, 0, 5015] - x[0, 0, 5000]) < 0.35
    pm = p.copy(); pm[:, 4] = 1                                           # Monitor = Delta: output - delayed input
    yd, _ = port.render(pm, xc)
    assert np.allclose(yd[0, 0, 5015], yc[0, 0, 5015] - xc[0, 0, 5000], atol=1e-6)


def test_var_amount_zero_is_transparent_and_more_amount_adds_hf():
    port = zo.PortPlugin("VAR")
    ns, nf = 1, 48000
    x = wl.programme_bursts(ns, nf)
    p = defaults(port, ns)
    p[:, 0] = 0.0
    y, _ = port.render(p, x)
    assert np.abs(y - x).max() < 1e-7                                     # g = 1, airGain = 0 (:160,176)
    p[:, 0] = 100.0
    y, _ = port.render(p, x)
    d = y - x
    assert np.abs(d).max() > 1e-4
    spec = np.abs(np.fft.rfft(d[0, 0]))
    f = np.fft.rfftfreq(nf, 1 / 48000.0)
    assert spec[f > 8000].sum() > 20 * spec[f < 2000].sum()               # what is added is high-pass / 16 kHz band-pass material


def test_red_ducks_the_wet_signal_only_when_the_reference_is_quiet():
    port = zo.PortPlugin("RED")
    nf = 48000
    x = wl.programme_bursts(1, nf, n_ch=6)
    x[:, 0:2] = 0.3 * np.sign(x[:, 0:2]) * np.abs(x[:, 0:2]) ** 0.3        # loud, dense wet signal
    p = defaults(port, 1)
    loud_ref = x.copy()
    loud_ref[:, 4:6] = x[:, 0:2]                                            # reference as loud as the wet: ratio of return 0 dB < threshold
    y, _ = port.render(p, loud_ref)
    assert np.array_equal(y[0, 2:], loud_ref[0, 2:])                        # ch3, ch4, ref pass through (:43)
    assert np.abs(y[0, :2] - loud_ref[0, :2]).max() < 1e-6
    quiet = x.copy()
    quiet[:, 4:6] *= 0.001                                                  # reference 60 dB down: the tail is ducked
    y, _ = port.render(p, quiet)
    g = np.abs(y[0, 0, 24000:]).mean() / np.abs(quiet[0, 0, 24000:]).mean()
    assert 10 ** (-12.5 / 20) < g < 0.9                                     # at most Amount = 12 dB
    assert np.array_equal(y[0, 2:], quiet[0, 2:])


@pytest.mark.parametrize("key", ["RED", "ClickBeGoneSG", "VAR", "GTS"])
def test_block_start_parameter_changes_and_threads(key):
    """Zones are written once per compute() call: a change takes effect at its host block's first frame; the multithreaded
    driver equals the single-threaded one."""
    port = zo.PortPlugin(key)
    ns, nf, block = 5, 4096, 512
    x = wl.programme_bursts(ns, nf, n_ch=port.n_in)
    if key == "RED":
        x[:, 4:6] *= 0.001                                                  # a quiet reference, or the leaf is a pass-through
        x[:, 0:2] = 0.3 * np.sign(x[:, 0:2]) * np.abs(x[:, 0:2]) ** 0.3
    p = wl.automation(port.sliders, nf // block, ns, which=[0, 1])
    y1, _ = port.render(p, x, block=block, threads=1)
    y4, _ = port.render(p, x, block=block, threads=4)
    assert np.array_equal(y1, y4)
    ps = np.repeat(p[:1], nf // block, axis=0)
    ys, _ = port.render(ps, x, block=block, threads=1)
    assert np.array_equal(ys[:, :, :block], y1[:, :, :block])               # block 0 has the same zones in both renders
    assert not np.array_equal(ys, y1)
