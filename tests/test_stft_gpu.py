"""GPU parity tests for the STFT class (plugin stft_mask2048, BASELINE config 4): the CUDA path through the
C ABI against the reference's own compiled DSP + its WDL FFT on the same seeded inputs.

Criteria (BASELINE.json north_star / SURVEY.md section 8c):
  fp64 mode: |gpu - ref| <= 1e-12 * max(|ref|, block peak) on the doubles before the float store;
  fp32 mode: residual vs the reference render below -100 dBFS RMS (re full scale 1.0);
  hop counter / ring index / frame count: exact."""
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
def ref():
    return zo.RefPlugin("stft_mask2048")


def make_case(ref, ns, nf, automate, block=512):
    x = wl.programme_bursts(ns, nf)
    base = np.array([s["default"] for s in ref.sliders], np.float32)
    if automate:
        p = wl.automation(ref.sliders, (nf + block - 1) // block, ns, which=[0, 1, 4], base=base)   # amount, thresh, mix
    else:
        p = np.tile(base, (ns, 1))
    return p, x


@pytest.mark.parametrize("automate,ns,nf", [(False, 5, 48000 + 13), (True, 6, 48000 + 700)])
def test_fp64_parity_and_counters(gpu, ref, automate, ns, nf):
    p, x = make_case(ref, ns, nf, automate)
    y, yd, v = ref.render(p, x, want_f64=True, want_vars=True)
    bp = gpu.BatchProcessor("stft_mask2048", ns, precision="f64").prepare(p if p.ndim == 2 else p[0])
    tap = np.zeros_like(yd)
    out = bp.process(x, tap64=tap, params=p if p.ndim == 3 else None)
    e = rel_err(tap, yd).max()
    assert e <= TOL64, e
    assert np.abs(out.astype(np.float64) - y).max() <= 2.0 ** -23 * np.abs(y).max()
    for name in ("hop_ctr", "ringpos", "frames_done"):          # integer / indexing state: exact
        assert np.array_equal(bp.read_state(name), v[name]), name
    bp.close()


def test_split_calls_with_ragged_sizes(gpu, ref):
    """State (rings, accumulators, per-bin EMAs, hop phase) carries across calls of any length.  Calls are
    cut on host-block boundaries of the one-shot render so both see the same parameter rows."""
    ns, nf = 3, 512 * 60 + 300
    p, x = make_case(ref, ns, nf, True)
    y, yd = ref.render(p, x, want_f64=True)
    bp = gpu.BatchProcessor("stft_mask2048", ns, precision="f64").prepare(p[0])
    tap = np.zeros_like(yd)
    for a, b in ((0, 512), (512, 512 * 7), (512 * 7, 512 * 31), (512 * 31, nf)):
        t = np.zeros((ns, 2, b - a))
        bp.process(np.ascontiguousarray(x[:, :, a:b]), tap64=t, params=p[a // 512:])
        tap[:, :, a:b] = t
    assert rel_err(tap, yd).max() <= TOL64
    # static parameters: any cut positions (hop phase not aligned to the calls)
    p0, x0 = make_case(ref, ns, 20000, False)
    y0, yd0 = ref.render(p0, x0, want_f64=True)
    bp.prepare(p0)
    tap0 = np.zeros_like(yd0)
    cuts = [0, 1, 100, 511, 513, 2048 + 77, 9999, 20000]
    for a, b in zip(cuts[:-1], cuts[1:]):
        t = np.zeros((ns, 2, b - a))
        bp.process(np.ascontiguousarray(x0[:, :, a:b]), tap64=t)
        tap0[:, :, a:b] = t
    assert rel_err(tap0, yd0).max() <= TOL64
    bp.close()


def test_identity_at_zero_amount(gpu, ref):
    """Size-independent property: with amount = 0 the masker is a pure 2048-frame delay (analysis x synthesis
    windows overlap-add to one), at any batch size."""
    ns, nf = 64, 48000
    p, x = make_case(ref, ns, nf, False)
    p[:, 0] = 0
    bp = gpu.BatchProcessor("stft_mask2048", ns, precision="f64").prepare(p)
    tap = np.zeros((ns, 2, nf))
    bp.process(x, tap64=tap)
    assert np.abs(tap[:, :, 2048:] - x[:, :, :-2048]).max() <= 1e-14
    assert np.abs(tap[:, :, :2048]).max() <= 1e-15       # only window-edge leakage of the first partial frames
    bp.close()


def test_fp32_mode_residual(gpu, ref):
    ns, nf = 8, 48000
    p, x = make_case(ref, ns, nf, True)
    y, yd = ref.render(p, x, want_f64=True)
    bp = gpu.BatchProcessor("stft_mask2048", ns, precision="f32").prepare(p[0])
    out = bp.process(x, params=p)
    r = out.astype(np.float64) - yd
    per_stream = np.array([dbfs_rms(r[s]) for s in range(ns)])
    print("fp32 residual dBFS rms per stream: min %.1f median %.1f max %.1f" % (per_stream.min(), np.median(per_stream), per_stream.max()))
    assert per_stream.max() < -100.0, per_stream.max()
    bp.close()


def test_matches_golden_from_reference(gpu):
    g = np.load(GOLD / "stft_mask2048_auto.npz")
    ns, nf, block = int(g["n_streams"]), int(g["n_frames"]), int(g["block"])
    x = wl.programme_bursts(ns, nf)
    p = g["params"]
    bp = gpu.BatchProcessor("stft_mask2048", ns, block_len=block, precision="f64").prepare(p[0])
    tap = np.zeros((ns, 2, nf))
    bp.process(x, tap64=tap, params=p)
    n = g["y64_head"].shape[-1]
    pk = np.abs(g["y64_tail"]).max(axis=2, keepdims=True)
    for got, want in ((tap[:, :, :n], g["y64_head"]), (tap[:, :, -n:], g["y64_tail"])):
        assert (np.abs(got - want) <= TOL64 * np.maximum(np.abs(want), pk)).all()
    assert np.allclose(np.abs(tap).sum(axis=2), g["y64_abs_sum"], rtol=1e-11, atol=0)
    bp.close()


def test_device_buffers_cfg4_shape(gpu, ref):
    """512 stereo streams x 2 s from device buffers; 6 streams re-rendered by the reference."""
    torch = pytest.importorskip("torch")
    ns, nf = 512, 96000
    sel = np.random.default_rng(11).choice(ns, 6, replace=False)
    xs = wl.programme_bursts(6, nf, first_stream=900)
    x = torch.zeros((ns, 2, nf), dtype=torch.float32, device="cuda")
    x[:] = torch.from_numpy(xs[0]).cuda()
    for i, s in enumerate(sel):
        x[int(s)] = torch.from_numpy(xs[i]).cuda()
    p = np.tile(np.array([s["default"] for s in ref.sliders], np.float32), (ns, 1))
    bp = gpu.BatchProcessor("stft_mask2048", ns, precision="f64").prepare(p)
    torch.cuda.synchronize()
    y = bp.process(x)
    bp.sync()
    yr, _ = ref.render(p[sel], xs)
    got = y[torch.from_numpy(sel).cuda()].cpu().numpy()
    assert np.abs(got.astype(np.float64) - yr).max() <= 2.0 ** -22 * max(np.abs(yr).max(), 1e-3)
    bp.close()
