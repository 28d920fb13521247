"""CPU tests of the Faust-leaf restatement (oracle/za_faust_port.c, PARITY UNPINNED -- no Faust compiler and no
reference-side render of a Faust leaf exists) against an independent numpy float32 transcription
of ModTilt.dsp, plus the properties the .dsp implies."""
import numpy as np

import za_oracle as zo
from za_gpu import workloads as wl

F = np.float32


def numpy_modtilt(p, x, srate=48000.0, coefs=None):
    """ModTilt.dsp:25-102 frame by frame in numpy float32 scalars (one stream, static zone values)."""
    tilt, pivot, mix = (F(v) for v in p)
    fs = F(min(192000.0, max(1.0, float(int(srate)))))
    pi = F(np.pi)

    def a_from_hz(hz):
        return F(1.0) - np.exp(((F(-2.0) * pi) * max(F(hz), F(0.001))) / fs, dtype=F)

    a_env, a_base, a_piv, a_trim = a_from_hz(25.0), a_from_hz(1.0), a_from_hz(pivot), a_from_hz(0.2)
    g_hi = np.power(F(10.0), (tilt * F(0.5)) / F(20.0), dtype=F)
    g_lo = np.power(F(10.0), ((-tilt) * F(0.5)) / F(20.0), dtype=F)
    if coefs is not None:      # slider-rate values of the C restatement: isolates the per-frame recursion from libm's last bit
        a_env, a_base, a_trim, a_piv, g_hi, g_lo = (F(v) for v in coefs[:6])
    env2 = base = lp = rs = mean = F(0.0)
    one = F(1.0)
    out = np.empty_like(x)
    for i in range(x.shape[1]):
        xl, xr = x[0, i], x[1, i]
        m_in = F(0.5) * (xl + xr)
        env2 = (m_in * m_in) * a_env + env2 * (one - a_env)
        env = np.sqrt(max(env2, F(0.0)))
        base = env * a_base + base * (one - a_base)
        m = env - base
        lp = m * a_piv + lp * (one - a_piv)
        m2_tilt = lp * g_lo + (m - lp) * g_hi
        m2 = m * F(0.25) + m2_tilt * F(0.75)
        env_tp = max(base + m2, F(0.05) * env)
        r0 = (env_tp + F(1e-9)) / (env + F(1e-9))
        r0c = min(max(r0, F(0.67)), F(1.5))
        rs = (r0c - one) * F(0.05) + rs * (one - F(0.05))
        g = one + rs
        rdb = F(20.0) * np.log10(max(g, F(1e-12)), dtype=F)
        mean = rdb * a_trim + mean * (one - a_trim)
        trim = np.power(F(10.0), (-mean) / F(20.0), dtype=F)
        out[0, i] = (xl * (one - mix) + (xl * g) * mix) * trim
        out[1, i] = (xr * (one - mix) + (xr * g) * mix) * trim
    return out


def test_port_matches_numpy_transcription():
    port = zo.PortPlugin("ModTilt")
    nf = 6000
    x = wl.programme_bursts(2, nf)
    p = np.array([[-6.0, 2.5, 1.0], [3.0, 5.0, 0.5]], np.float32)
    y, _ = port.render(p, x, threads=1)
    for s in range(2):
        # fully independent: numpy's exp / pow may differ from libm's in the last bit, and 1 - exp(-w) for the slow
        # smoothers (w ~ 2.6e-5) turns one ulp of exp into 2e-3 of the coefficient
        want = numpy_modtilt(p[s], x[s])
        assert np.abs(y[s] - want).max() <= 1e-4 * max(np.abs(want).max(), 1e-3), s
        # same slider-rate values: the per-frame chain agrees to the last bits of log10 / pow
        want = numpy_modtilt(p[s], x[s], coefs=port.coefs(48000.0, *p[s]))
        assert np.abs(y[s] - want).max() <= 4e-7 * max(np.abs(want).max(), 1e-3), s


def test_port_properties():
    port = zo.PortPlugin("ModTilt")
    ns, nf = 6, 20000
    x = wl.programme_bursts(ns, nf)
    # tilt = 0: both band gains are 1 -> ratio 1 -> pass-through to float rounding
    y, _ = port.render(np.tile(np.array([0.0, 3.0, 1.0], np.float32), (ns, 1)), x)
    assert np.abs(y - x).max() <= 1e-5
    # the gain is linked: L and R of a stream see the same factor
    y, _ = port.render(np.tile(np.array([-6.0, 3.0, 1.0], np.float32), (ns, 1)), x)
    big = (np.abs(x[:, 0]) > 1e-3) & (np.abs(x[:, 1]) > 1e-3)
    gl, gr = y[:, 0][big] / x[:, 0][big], y[:, 1][big] / x[:, 1][big]
    assert np.abs(gl - gr).max() < 1e-5
    assert 0.6 < gl.min() and gl.max() < 1.7 and np.ptp(gl) > 1e-3          # clamp range :85, and the effect does something
    # zone values land on the declared grid and inside the declared range (FaustJuceProcessor.cpp:379)
    a, _ = port.render(np.tile(np.array([-6.04, 3.004, 1.7], np.float32), (ns, 1)), x)
    b, _ = port.render(np.tile(np.array([-6.0, 3.0, 1.0], np.float32), (ns, 1)), x)
    assert np.array_equal(a, b)
    # threads and host-block length do not change static-parameter results; per-block rows apply at block starts
    c, _ = port.render(np.tile(np.array([-6.0, 3.0, 1.0], np.float32), (ns, 1)), x, block=37, threads=3)
    assert np.array_equal(b, c)
    nb = (nf + 99) // 100
    p = np.tile(np.array([-6.0, 3.0, 1.0], np.float32), (nb, ns, 1)); p[50:, :, 2] = 0.0
    d, _ = port.render(p, x, block=100)
    assert np.array_equal(d[:, :, :5000], b[:, :, :5000]) and not np.array_equal(d[:, :, 5000:5100], b[:, :, 5000:5100])


def test_library_lists_the_faust_leaf():
    import za_gpu
    assert "ModTilt" in za_gpu.plugin_keys()
    info = za_gpu.plugin_info("ModTilt")
    assert (info.n_in, info.n_out, info.kind, info.n_sliders) == (2, 2, 1, 3)
    lay = za_gpu.slider_layout("ModTilt")
    ref = zo.parse_sliders(zo.plugin_source_text("ModTilt"))
    assert [float(s.def_) for s in lay] == [s["default"] for s in ref] == [0.0, 3.0, 1.0]


def test_kernel_arithmetic_on_host_equals_restatement(tmp_path):
    """csrc/modtilt_core.h (the CUDA kernel's per-frame code) compiled by g++ with libm in place of the MUFU
    approximations: same expression order as the restatement -> the same float output, except where the branch-free
    square root rounds differently from sqrtf (a 1-ulp case about once in 10^7 operands)."""
    import ctypes as C
    import subprocess
    from pathlib import Path
    repo = Path(__file__).resolve().parent.parent
    so = tmp_path / "libemu_modtilt.so"
    subprocess.run(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-mfma", "-shared", "-fPIC",
                    "-I" + str(repo / "zorakaudio-experimental-plugins_b200" / "csrc"), "-I/usr/local/cuda/include",
                    str(repo / "tools" / "emu_modtilt.cpp"), "-o", str(so)], check=True)
    emu = C.CDLL(str(so))
    port = zo.PortPlugin("ModTilt")
    ns, nf = 6, 48000 + 5
    x = wl.programme_bursts(ns, nf)
    p = np.array([[-6.0, 2.0, 1.0], [3.0, 5.0, 0.5], [-4.5, 3.0, 0.8], [-0.1, 2.01, 0.001], [1.5, 4.2, 0.25], [-6.0, 5.0, 0.0]], np.float32)
    y, _ = port.render(p, x)
    coef = np.stack([port.coefs(48000.0, *p[s]) for s in range(ns)]).astype(np.float32)
    out = np.zeros_like(y)
    emu.emu_modtilt_render(C.c_void_p(coef.ctypes.data), C.c_void_p(x.ctypes.data), C.c_void_p(out.ctypes.data), ns, C.c_long(nf), 0)
    # the stage-by-stage 16-frame form the kernel runs inside a host block is the same arithmetic
    out16 = np.zeros_like(y)
    emu.emu_modtilt_render(C.c_void_p(coef.ctypes.data), C.c_void_p(x.ctypes.data), C.c_void_p(out16.ctypes.data), ns, C.c_long(nf), 1)
    assert np.array_equal(out16, out)
    assert np.mean(out == y) > 0.999
    assert np.abs(out.astype(np.float64) - y).max() <= 1e-6 * np.abs(y).max()


import ctypes as C  # noqa: E402


def _build_emu_scan(tmp_path):
    import subprocess
    from pathlib import Path
    repo = Path(__file__).resolve().parent.parent
    so = tmp_path / "libemu_scan.so"
    subprocess.run(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-mfma", "-shared", "-fPIC",
                    "-I" + str(repo / "zorakaudio-experimental-plugins_b200" / "csrc"), "-I/usr/local/cuda/include",
                    str(repo / "tools" / "emu_scan.cpp"), "-o", str(so)], check=True)
    return C.CDLL(str(so))


def test_scan_kernel_tile_code_on_host_nulls_against_the_restatement(tmp_path):
    """The time-parallel ModTilt kernel's tile body (csrc/modtilt_scan_core.h: pass A, 32-lane Kogge-Stone, pass B, exact
    tile carries) compiled for the host with lanes.h's 32-float warp type: 6 s with the tilt AND the pivot moving every
    512-frame block must null against the sequential float restatement far below the -100 dBFS criterion, and the five
    smoother states at the end must agree to float rounding."""
    emu = _build_emu_scan(tmp_path)
    port = zo.PortPlugin("ModTilt")
    block, nb = 512, 562
    nf = block * nb
    for stream, level in ((0, 1.0), (1, 0.05)):
        x = wl.programme_bursts(1, nf, first_stream=stream) * np.float32(level)
        rng = np.random.default_rng(stream)
        p = np.empty((nb, 1, 3), np.float32)
        p[:, 0, 0] = np.round(np.clip(-1.5 + 4.0 * np.sin(np.arange(nb) / 17.0) + rng.normal(0, 0.3, nb), -6, 3), 1)
        p[:, 0, 1] = np.round(np.clip(3.5 + 1.4 * np.sin(np.arange(nb) / 29.0), 2, 5), 2)
        p[:, 0, 2] = 0.8
        y, _ = port.render(p, x, block=block)
        coef = np.zeros((nb, 8), np.float32)
        for b in range(nb):
            c = port.coefs(48000.0, *p[b, 0])            # a_env a_base a_trim a_piv g_hi g_lo mix fs
            coef[b, :6] = [c[3], np.float32(1.0) - c[3], c[5], c[4], c[6], np.float32(1.0) - c[6]]
        fixed3 = np.array(c[:3], np.float32)
        for fast in (0, 1):                  # correctly rounded maps / the kernel's throughput forms (host: libm stand-ins)
            state = np.zeros(5, np.float32)
            out = np.zeros_like(x)
            # two calls: state carries across calls exactly like across tiles
            cut = block * 200
            for a, b_ in ((0, cut), (cut, nf)):
                emu.emu_modtilt_scan(C.c_void_p(coef[a // block:].ctypes.data), nb - a // block, block, C.c_void_p(fixed3.ctypes.data),
                                     C.c_void_p(state.ctypes.data), C.c_void_p(x[0, 0, a:].ctypes.data), C.c_void_p(out[0, 0, a:].ctypes.data),
                                     C.c_long(b_ - a), C.c_long(nf), fast)
            r = out.astype(np.float64) - y
            db = 20 * np.log10(np.sqrt(np.mean(r ** 2)) + 1e-300)
            rel = np.abs(r).max() / np.abs(y).max()
            print(f"scan emulation vs restatement (level {level}, fast {fast}): residual {db:.1f} dBFS rms, max |d| / peak {rel:.2e}, equal {np.mean(out == y):.3f}")
            assert db < -120.0, db
            assert rel < 2e-5
