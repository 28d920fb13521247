"""CPU tests of the DSP-JSFX -> CUDA translator (za_gpu/jsfxc.py): the generated translation units compiled as
plain C++ (tools/emu_jsfx.py, -DZA_JSFX_HOST_EMU: same generated code, host stand-ins for the device intrinsics)
must reproduce the reference's compiled DSP (oracle/_ref) bit for bit, final variables included."""
import sys
from pathlib import Path

import pytest

REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO / "tools"))
import za_oracle as zo


@pytest.mark.parametrize("key,auto", [("RTT", True), ("SOMA", False), ("ATTACK", True), ("ERBTilt", False),
                                      ("SpectralStabilizer", True), ("DDT", False), ("DPT", True), ("SaliencePush", False),
                                      ("ADS", True), ("BedRock", False), ("Roomalizer", True), ("TSEQ", False),
                                      ("semprobe", True), ("Alias", True), ("Alias", False)])
def test_translation_is_bit_identical_on_host(key, auto):
    if not zo.have_ref(key):
        pytest.skip("oracle/_ref not built")
    import emu_jsfx
    err, bad = emu_jsfx.compare(key, nf=9000, automate=auto, verbose=False)
    assert err == 0.0, err
    assert not bad, bad[:4]


def test_nan_conditions_and_ordered_compares_match_the_reference():
    """plugins/Workloads/semprobe: 0/0 and inf - inf reach conditions, `!=`, `&&`, `||`, `while`, min / max.  The reference
    compiles ordered compares (a NaN condition is false, `a != b` with a NaN operand is 0: dsp_jsfx_aot.py:3725, 4349-4352);
    the generated code must take the same branches -- outputs and every variable bit-identical, NaN payloads included."""
    import numpy as np
    import emu_jsfx
    from za_gpu import workloads as wl
    if not zo.have_ref("semprobe"):
        pytest.skip("oracle/_ref not built")
    ref = zo.RefPlugin("semprobe")
    e = emu_jsfx.Emu("semprobe")
    nf = 6000
    x = wl.programme_bursts(1, nf, n_ch=2)
    x[0, 0, 100], x[0, 1, 200], x[0, 0, 300] = np.inf, np.nan, -np.inf
    x[0, 0, 400:420] = 0.0
    for sel in range(4):
        p = np.array([[0.5, sel]], np.float32)
        y, yd, v = ref.render(p, x, want_f64=True, want_vars=True)
        ye, te, ve = e.render(ref.sliders, p, x[0])
        assert np.array_equal(te.view(np.uint64), yd[0].view(np.uint64)), sel
        for n in v:
            if n in ve:
                assert ve[n] == v[n][0] or (np.isnan(ve[n]) and np.isnan(v[n][0])), (sel, n)
        assert v["f2"][0] == 0.0 and v["k"][0] == 0.0 and v["w"][0] == 0.0       # the ordered forms really were exercised


@pytest.mark.parametrize("key,auto", [("decrust512", True), ("decrust512", False), ("stft_mask2048", True)])
def test_fft_hook_scripts_match_the_reference_and_wdl(key, auto):
    """Scripts that call the FFT runtime hooks: the translator's per-stream device transforms (csrc/jsfx_rt.cuh: Stockham
    radix-2 + scatter into WDL's order, the two-for-one real split, convolve_c) against the reference's compiled path
    over its own WDL FFT.  Not bit-identical (different butterfly order), far inside the 1e-12 criterion."""
    if not zo.have_ref(key):
        pytest.skip("oracle/_ref not built")
    import emu_jsfx
    err, bad = emu_jsfx.compare(key, nf=9000, automate=auto, verbose=False)
    assert err <= 1e-13, err
    assert not bad, bad[:4]


@pytest.mark.parametrize("log2n", [4, 5, 7, 10, 13, 15])
def test_every_fft_hook_at_every_size(log2n):
    """plugins/Workloads/fftprobe: fft (WDL's permuted order visible), fft + fft_permute (natural order), fft + ifft,
    fft_real, fft_real + ifft_real, permute + ipermute at sizes 16 .. 32768 against src/WDL/fft.c behind the reference's
    hooks (oracle/ref_fft_hooks.c)."""
    import numpy as np
    import emu_jsfx
    from za_gpu import workloads as wl
    if not zo.have_ref("fftprobe"):
        pytest.skip("oracle/_ref not built")
    ref = zo.RefPlugin("fftprobe")
    e = emu_jsfx.Emu("fftprobe")
    n = 1 << log2n
    nf = 3 * n + 5
    x = wl.programme_bursts(1, nf, n_ch=2)
    for mode in range(6):
        p = np.array([[log2n, mode]], np.float32)
        y, yd = ref.render(p, x, want_f64=True)
        ye, te, ve = e.render(ref.sliders, p, x[0])
        assert np.abs(yd).max() > 0
        d = np.abs(te - yd[0]).max() / np.abs(yd).max()
        assert d <= (0.0 if mode == 5 else 5e-15), (mode, d)


def test_parser_precedence_and_forms():
    from za_gpu import jsfxc
    P = lambda s: jsfxc.Parser(s).program()
    assert P("-x^2")[0] == ("bin", "^", ("un", "-", ("var", "x")), ("num", 2.0))          # prefix binds tighter than ^
    assert P("a ? b")[0] == ("tern", ("var", "a"), ("var", "b"), ("num", 0.0))            # missing ':' -> 0
    assert P("a = b = 1")[0][0] == "assign" and P("a = b = 1")[0][3][0] == "assign"       # right associative
    assert P("a\n  || b")[0] == ("bin", "||", ("var", "a"), ("var", "b"))                 # newline + infix continues
    assert len(P("a\n-b")) == 2                                                            # '-' may start a statement
    assert P("x & 3 == 3")[0][1] == "=="  or P("x & 3 == 3")[0][1] == "&"                 # same level, left to right
    assert P("loop(3, a += 1; b += 2)")[0][0] == "loop"
    # instance() / this.: a namespaced call makes the instance variable a member of the call-site pseudo-object
    u = jsfxc.Unit("t", "@init\nfunction f(x) instance(s) ( s += x; this.n += 1; s; );\n@sample\na.f(spl0); b.f(spl1); f(1);\n")
    assert {"a.s", "a.n", "b.s", "b.n", "f.s", "f.n"} <= set(u.vars)
    with pytest.raises(NotImplementedError):
        jsfxc.Unit("t", "@sample\nx = gmem[3];\n")


@pytest.mark.parametrize("key", ["RTT", "SOMA", "ATTACK", "ERBTilt", "SpectralStabilizer", "DDT", "DPT", "SaliencePush", "ADS",
                                 "BedRock", "Roomalizer", "TSEQ"])
def test_translation_matches_committed_golden(key):
    """Same check against tests/golden/translated_leaves.npz (reference outputs generated by tests/golden/make_golden.py),
    so it also runs where neither /root/reference nor oracle/_ref exists."""
    import hashlib
    import numpy as np
    import emu_jsfx
    from za_gpu import workloads as wl
    g = np.load(Path(__file__).parent / "golden" / "translated_leaves.npz")
    sliders = zo.parse_sliders(zo.plugin_source_text(key))
    e = emu_jsfx.Emu(key)
    nf = 9000
    base = np.array([s["default"] for s in sliders], np.float32)
    p = wl.automation(sliders, (nf + 511) // 512, 1, which=list(range(min(3, len(base)))), base=base)
    x = wl.programme_bursts(1, nf, n_ch=e.n_in)
    y, tap, _ = e.render(sliders, p[:, 0], x[0])
    assert np.array_equal(tap[:2, -1024:], g[key + "_y64_tail"])
    assert hashlib.sha256(tap[None].tobytes()).digest() == g[key + "_y64_sha256"].tobytes()


def test_cfg1_ddt_on_host():
    """BASELINE config 1 (DDT, 10 s x 2 ch, block 512, declared defaults, the survey's gated-noise + sine input): the
    generated translation unit, emulated on the host, is bit-identical to the reference's compiled path."""
    import numpy as np
    import emu_jsfx as emu
    from za_gpu import workloads as wl
    if not zo.have_ref("DDT"):
        pytest.skip("oracle/_ref not built")
    ref = zo.RefPlugin("DDT")
    x = wl.cfg1_input()
    p = np.array([[s["default"] for s in ref.sliders]], np.float32)
    y, yd, v = ref.render(p, x, want_f64=True, want_vars=True)
    ye, te, ve = emu.Emu("DDT").render(ref.sliders, p, x[0])
    assert np.array_equal(te, yd[0])
    assert np.array_equal(ye, y[0])
    for n in v:
        if n in ve:
            assert ve[n] == v[n][0], n


@pytest.mark.parametrize("f64", [True, False])
def test_heap_overflow_is_reported_not_silent(f64):
    """A mem[] index beyond the heap a leaf was sized for must fail the call (the reference would grow the heap; the
    batch renderer has a fixed one): the runtime clamps the access to its spare slot and raises the error flag."""
    import numpy as np
    import emu_jsfx as emu
    from za_gpu import workloads as wl
    if not zo.have_ref("SpectralStabilizer"):
        pytest.skip("oracle/_ref not built")
    ref = zo.RefPlugin("SpectralStabilizer")
    e = emu.Emu("SpectralStabilizer", f64=f64)
    x = wl.programme_bursts(1, 2048, n_ch=2)[0]
    p = np.array([s["default"] for s in ref.sliders], np.float32)
    e.render(ref.sliders, p, x)                                   # the real heap size: fine
    with pytest.raises(AssertionError, match="overflow"):
        e.render(ref.sliders, p, x, mem_slots=1000)              # the script keeps state up to slot ~2080
