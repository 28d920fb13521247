"""GPU parity tests for the linear-IIR class (plugin cascade8, BASELINE config 2): the CUDA path
through the C ABI against the oracle on the same seeded inputs.

Criteria (BASELINE.json north_star / SURVEY.md section 8c):
  fp64 mode: |gpu - ref| <= 1e-12 * max(|ref|, block peak)  on the doubles BEFORE the float store;
  fp32 mode: residual vs the reference render below -100 dBFS RMS (re full scale 1.0);
  state carry, block positions of parameter changes: exact behaviour of the reference host."""
from pathlib import Path

import numpy as np
import pytest

import za_oracle as zo
from za_gpu import workloads as wl
from parity import TOL64, block_peak, dbfs_rms, rel_err  # noqa: F401  (the one criterion definition, tests/parity.py)

pytestmark = pytest.mark.gpu
GOLD = Path(__file__).parent / "golden"


@pytest.fixture(scope="module")
def gpu(built_lib):
    import za_gpu
    if za_gpu.lib().za_gpu_device_count() < 1:
        pytest.fail("GPU test run without a CUDA device")
    return za_gpu


@pytest.fixture(scope="module")
def case():
    ns, nf = 48, 48000 * 2 + 13
    p, x = wl.cascade8_params(ns), wl.noise_two_sines(ns, nf)
    y, yd = zo.best_oracle("cascade8").render(p, x, want_f64=True)
    return p, x, y, yd


@pytest.mark.parametrize("warps", [0, 1, 2, 4])
@pytest.mark.parametrize("generic", [0, 1])
def test_fp64_parity(gpu, case, warps, generic):
    p, x, y, yd = case
    bp = gpu.BatchProcessor("cascade8", x.shape[0], precision="f64")
    bp.set_option("warps_per_stream", warps); bp.set_option("force_generic_io", generic)
    bp.prepare(p)
    tap = np.zeros_like(yd)
    out = bp.process(x, tap64=tap)
    e = rel_err(tap, yd).max()
    assert e <= TOL64, e
    # the float store is the reference's (float) cast: equal wherever the doubles round the same way
    assert np.mean(out == y) > 0.9999
    assert np.abs(out.astype(np.float64) - y).max() <= 2.0 ** -23 * np.abs(y).max()
    for i, name in enumerate(["z11", "lp8"]):
        assert np.isfinite(bp.read_state(name)).all()
    bp.close()


def test_fp64_matches_golden_from_reference(gpu):
    for tag in ("", "_auto"):
        g = np.load(GOLD / f"cascade8{tag}.npz")
        ns, nf, block = int(g["n_streams"]), int(g["n_frames"]), int(g["block"])
        x = wl.noise_two_sines(ns, nf)
        params = g["params"]
        bp = gpu.BatchProcessor("cascade8", ns, block_len=block, precision="f64")
        bp.prepare(params if params.ndim == 2 else params[0])
        tap = np.zeros((ns, 1, nf))
        bp.process(x, tap64=tap, params=params if params.ndim == 3 else None)
        n = g["y64_head"].shape[-1]
        for got, ref in ((tap[:, :, :n], g["y64_head"]), (tap[:, :, -n:], g["y64_tail"])):
            assert (np.abs(got - ref) <= TOL64 * np.maximum(np.abs(ref), np.abs(ref).max(axis=2, keepdims=True))).all()
        assert np.allclose(np.abs(tap).sum(axis=2), g["y64_abs_sum"], rtol=1e-11, atol=0)
        bp.close()


def test_state_carries_across_calls_and_ragged_sizes(gpu, case):
    p, x, y, yd = case
    ns, _, nf = x.shape
    bp = gpu.BatchProcessor("cascade8", ns, precision="f64").prepare(p)
    cuts = [0, 1, 32, 545, 4096 + 5, 50001, nf]        # block-misaligned, sub-chunk and multi-tile pieces
    tap = np.zeros_like(yd)
    for a, b in zip(cuts[:-1], cuts[1:]):
        xi = np.ascontiguousarray(x[:, :, a:b]); ti = np.zeros((ns, 1, b - a))
        bp.process(xi, tap64=ti)
        tap[:, :, a:b] = ti
    # the reference restarts its host-block grid at every processBlock call; cascade8 has no @block
    # dependence, so the split render must equal the one-shot render
    assert rel_err(tap, yd).max() <= TOL64
    z = bp.read_state("lp8")
    assert np.abs(z - yd[:, 0, -1]).max() <= TOL64 * np.abs(yd).max()      # lp8 is the last output sample
    bp.close()


def test_device_buffers_in_place_and_strided(gpu, case):
    torch = pytest.importorskip("torch")
    p, x, y, yd = case
    ns, _, nf = x.shape
    stride = (nf + 63) // 64 * 64 + 64
    buf = torch.zeros((ns, 1, stride), dtype=torch.float32, device="cuda")
    buf[:, :, :nf] = torch.from_numpy(x).cuda()
    bp = gpu.BatchProcessor("cascade8", ns, precision="f64").prepare(p)
    torch.cuda.synchronize()
    bp.process(buf, out=buf, n_frames=nf, frame_stride=stride)      # in == out
    bp.sync()
    got = buf[:, :, :nf].cpu().numpy()
    assert np.mean(got == y) > 0.9999
    assert (buf[:, :, nf:] == 0).all()                              # nothing written past n_frames
    bp.close()


def test_fp32_mode_residual(gpu, case):
    p, x, y, yd = case
    bp = gpu.BatchProcessor("cascade8", x.shape[0], precision="f32").prepare(p)
    out = bp.process(x)
    r = out.astype(np.float64) - yd
    per_stream = np.array([dbfs_rms(r[s]) for s in range(x.shape[0])])
    print("fp32 residual dBFS rms per stream: min %.1f median %.1f max %.1f" % (per_stream.min(), np.median(per_stream), per_stream.max()))
    assert per_stream.max() < -100.0, per_stream.max()
    bp.close()


@pytest.mark.parametrize("warps,nbuf,scalar", [(1, 1, 0), (1, 2, 0), (2, 1, 0), (2, 2, 0), (4, 1, 0), (4, 2, 0), (2, 0, 1)])
def test_fp32_kernel_variants(gpu, case, warps, nbuf, scalar):
    """Every launch shape of the packed-FP32 (FFMA2) kernel and the scalar fp32 kernel: residual vs the
    fp64 reference render below -100 dBFS per stream, and a split render (state carried through HBM,
    tails on the generic kernel) that agrees with the one-shot render to fp32 noise."""
    p, x, y, yd = case
    ns, _, nf = x.shape
    bp = gpu.BatchProcessor("cascade8", ns, precision="f32")
    bp.set_option("warps_per_stream", warps); bp.set_option("tile_buffers", nbuf); bp.set_option("scalar_f32", scalar)
    bp.prepare(p)
    out = bp.process(x)
    r = out.astype(np.float64) - yd
    per_stream = np.array([dbfs_rms(r[s]) for s in range(ns)])
    assert per_stream.max() < -100.0, per_stream.max()
    bp.prepare(p)
    cuts = [0, 64, 64 + 4096 * warps, 50016, nf]
    out2 = np.empty_like(out)
    for a, b in zip(cuts[:-1], cuts[1:]):
        out2[:, :, a:b] = bp.process(np.ascontiguousarray(x[:, :, a:b]))
    assert dbfs_rms(out2.astype(np.float64) - yd) < -100.0
    assert dbfs_rms(out2.astype(np.float64) - out) < -100.0
    bp.close()


def test_empty_and_argument_errors(gpu):
    import ctypes as C
    bp = gpu.BatchProcessor("cascade8", 3, precision="f64")
    x = np.zeros((3, 1, 16), np.float32)
    with pytest.raises(gpu.ZaError):
        bp.process(x)                                   # process before prepare
    bp.prepare(None)                                    # declared defaults
    out = bp.process(np.zeros((3, 1, 0), np.float32))   # empty input is legal
    assert out.shape == (3, 1, 0)
    with pytest.raises(gpu.ZaError):
        bp.process(x, tap64=None, frame_stride=8)       # stride < n_frames
    bp32 = gpu.BatchProcessor("cascade8", 3, precision="f32").prepare(None)
    with pytest.raises(gpu.ZaError):
        bp32.process(x, tap64=np.zeros((3, 1, 16)))     # tap needs fp64
    bp.close(); bp32.close()


def test_large_batch_spot_checked_against_oracle(gpu):
    """1024 streams x 10 s rendered from device buffers; 12 random streams re-rendered by the oracle."""
    torch = pytest.importorskip("torch")
    ns, nf = 1024, 480000
    p = wl.cascade8_params(ns)
    g = torch.Generator(device="cuda").manual_seed(1234)
    x = (torch.rand((ns, 1, nf), device="cuda", generator=g) - 0.5) * 0.5
    t = torch.arange(nf, device="cuda", dtype=torch.float32) / 48000.0
    x += 0.125 * torch.sin(2 * np.pi * 220.0 * t) + 0.125 * torch.sin(2 * np.pi * 3000.0 * t)
    bp = gpu.BatchProcessor("cascade8", ns, precision="f64").prepare(p)
    torch.cuda.synchronize()
    y = bp.process(x)
    bp.sync()
    pick = np.random.default_rng(5).choice(ns, 12, replace=False)
    xs = x[pick].cpu().numpy()
    yr, _ = zo.best_oracle("cascade8").render(p[pick], xs)
    got = y[pick].cpu().numpy()
    assert np.mean(got == yr) > 0.9999
    assert np.abs(got.astype(np.float64) - yr).max() <= 2.0 ** -23 * np.abs(yr).max()
    # linearity (size-independent property): render(2x) == 2 render(x) exactly (power-of-two scaling)
    bp.prepare(p)
    y2 = bp.process(x * 2.0)
    bp.sync()
    assert torch.equal(y2, y * 2.0)
    bp.close()


def _auto_case(ns, nf, block, rows=None, which=None):
    ref = zo.best_oracle("cascade8")
    x = wl.noise_two_sines(ns, nf)
    nb = (nf + block - 1) // block if rows is None else rows
    base = wl.cascade8_params(ns)
    p = wl.automation(ref.sliders, nb, ns, which=list(range(15)) if which is None else which, base=base[0])
    p[:, :, :] = np.where(np.isfinite(p), p, base[None])
    return ref, p, x


@pytest.mark.parametrize("ns,nf,block,rows", [(37, 512 * 40 + 77, 512, None), (9, 100 * 61 + 3, 100, None), (5, 512 * 12, 512, 5), (3, 31 * 7 + 5, 31, None)])
def test_automated_render_in_one_launch_per_segment(gpu, ns, nf, block, rows):
    """Per-block automation of all 15 sliders through cascade8_auto_kernel (scan tables rebuilt on the device per host block):
    fp64 within 1e-12 of level of the reference render and the filter state equal to rounding, for block lengths that do and
    do not divide the tiles, a ragged end, fewer parameter rows than host blocks (the last row persists); not one launch per
    block; and the same bits as the old launch-per-block path (option force_generic_io) to the criterion."""
    ref, p, x = _auto_case(ns, nf, block, rows)
    y, yd, v = ref.render(p, x, want_f64=True, want_vars=True, block=block)
    bp = gpu.BatchProcessor("cascade8", ns, 48000.0, block, "f64").prepare(p[0])
    tap = np.zeros_like(yd)
    l0 = bp.launch_count()
    out = bp.process(x, tap64=tap, params=p)
    nbq = min(len(p), (nf + block - 1) // block)
    assert bp.launch_count() - l0 <= 2 * ((nbq + 255) // 256) + 4, bp.launch_count() - l0
    assert rel_err(tap, yd).max() <= TOL64
    assert np.mean(out == y) > 0.9999
    for name in ("z11", "z32", "z62", "lp7", "lp8"):
        assert np.allclose(bp.read_state(name), v[name], rtol=1e-9, atol=1e-14), name
    bp.close()
    bp = gpu.BatchProcessor("cascade8", ns, 48000.0, block, "f64").prepare(p[0])
    bp.set_option("force_generic_io", 1)                 # the block-by-block path of the static kernels
    tap2 = np.zeros_like(yd)
    bp.process(x, tap64=tap2, params=p)
    assert rel_err(tap2, yd).max() <= TOL64 and rel_err(tap, tap2).max() <= TOL64
    bp.close()


def test_automated_render_fp32_and_what_follows_it(gpu):
    """Float mode of the automation kernel (< -100 dBFS per stream); a static call after an automated one continues from its
    state with the LAST row's coefficients (the static kernels' tables follow); an automated render fed in two calls gives
    the bits of one call."""
    ns, block = 21, 512
    nf = block * 30
    ref, p, x = _auto_case(ns, nf, block)
    xs = wl.noise_two_sines(ns, 4096 + 11, first_stream=100)
    pall = np.concatenate([p, np.repeat(p[-1:], (xs.shape[-1] + block - 1) // block, axis=0)], axis=0)
    _, yd = ref.render(pall, np.concatenate([x, xs], axis=2), want_f64=True, block=block)
    for prec in ("f32", "f64"):
        bp = gpu.BatchProcessor("cascade8", ns, 48000.0, block, prec).prepare(p[0])
        a = bp.process(x, params=p)
        b = bp.process(xs)                                # static: no parameter rows
        got = np.concatenate([a, b], axis=2).astype(np.float64)
        worst = max(dbfs_rms(got[s] - yd[s]) for s in range(ns))
        print(f"cascade8 automated then static, {prec}: worst stream {worst:.1f} dBFS")
        assert worst < (-100.0 if prec == "f32" else -120.0), worst   # f64: what is left is the float store of a hot (resonant, Q up to 8) signal
        bp.close()
        bp = gpu.BatchProcessor("cascade8", ns, 48000.0, block, prec).prepare(p[0])
        cut = block * 11
        parts = [bp.process(np.ascontiguousarray(x[:, :, :cut]), params=np.ascontiguousarray(p[:11])),
                 bp.process(np.ascontiguousarray(x[:, :, cut:]), params=np.ascontiguousarray(p[11:]))]
        assert np.array_equal(np.concatenate(parts, axis=2), a), prec
        bp.close()


def test_automated_render_longer_than_a_segment(gpu):
    """More host blocks than one 256-block segment of coefficient rows, with sliders that move only in some blocks (the
    unchanged rows are copies and the kernel keeps its tables across them)."""
    ns, block = 4, 64
    nf = block * 700 + 9
    ref, p, x = _auto_case(ns, nf, block, which=[0, 4, 14])
    p[100:400] = p[100]                                    # a long stretch without movement, across the segment boundary
    _, yd = ref.render(p, x, want_f64=True, block=block)
    bp = gpu.BatchProcessor("cascade8", ns, 48000.0, block, "f64").prepare(p[0])
    tap = np.zeros_like(yd)
    bp.process(x, tap64=tap, params=p)
    assert rel_err(tap, yd).max() <= TOL64
    bp.close()
