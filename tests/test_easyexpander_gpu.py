"""GPU parity tests for the nonlinear dynamics class (plugin EasyExpander, BASELINE config 3): the CUDA
path through the C ABI against the reference's own compiled DSP on the same seeded inputs.

Criteria (BASELINE.json north_star / SURVEY.md section 8c):
  fp64 mode: |gpu - ref| <= 1e-12 * max(|ref|, block peak) on the doubles before the float store;
  fp32 mode: residual vs the reference render below -100 dBFS RMS (re full scale 1.0);
  integer state (the hysteresis flag `expanding`) and the host-block positions of parameter changes: exact."""
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
    return zo.RefPlugin("EasyExpander")


def make_case(ref, ns, nf, automate, detector_filters=False, block=512):
    x = wl.programme_bursts(ns, nf)
    nb = (nf + block - 1) // block
    base = np.array([s["default"] for s in ref.sliders], np.float32)
    if detector_filters:
        base[3], base[4] = 120.0, 6000.0
    if automate:
        p = wl.automation(ref.sliders, nb, ns, which=[0, 1], base=base)     # threshold and depth move every block
    else:
        p = np.tile(base, (ns, 1))
    return p, x


@pytest.mark.parametrize("automate,filters,ns,nf", [(False, False, 21, 48000 + 13), (True, False, 16, 48000 * 2 + 5),
                                                    (True, True, 7, 30000)])
def test_fp64_parity_and_state(gpu, ref, automate, filters, ns, nf):
    p, x = make_case(ref, ns, nf, automate, filters)
    y, yd, v = ref.render(p, x, want_f64=True, want_vars=True)
    bp = gpu.BatchProcessor("EasyExpander", ns, precision="f64").prepare(p if p.ndim == 2 else p[0])
    tap = np.zeros_like(yd)
    out = bp.process(x, tap64=tap, params=p if p.ndim == 3 else None)
    e = rel_err(tap, yd).max()
    assert e <= TOL64, e
    assert np.mean(out == y) > 0.9999
    assert np.array_equal(bp.read_state("expanding"), v["expanding"])        # integer state: exact
    for name in ("gain_db_s", "det_slow", "det_db_m", "gr_db_m", "hpf_lp", "lpf_y"):
        got, want = bp.read_state(name), v[name]
        assert np.abs(got - want).max() <= 1e-11 * max(1.0, np.abs(want).max()), name
    bp.close()


def test_split_calls_match_one_shot(gpu, ref):
    """State carries across calls; a call boundary on a host-block boundary keeps the block grid."""
    ns, nf = 9, 512 * 40 + 77
    p, x = make_case(ref, ns, nf, True)
    y, yd = ref.render(p, x, want_f64=True)
    bp = gpu.BatchProcessor("EasyExpander", ns, precision="f64").prepare(p[0])
    tap = np.zeros_like(yd)
    for a, b in ((0, 512 * 3), (512 * 3, 512 * 17), (512 * 17, nf)):
        t = np.zeros((ns, 2, b - a))
        bp.process(np.ascontiguousarray(x[:, :, a:b]), tap64=t, params=p[a // 512:])
        tap[:, :, a:b] = t
    assert rel_err(tap, yd).max() <= TOL64
    bp.close()


def test_fp32_mode_residual(gpu, ref):
    ns, nf = 16, 48000 * 2
    p, x = make_case(ref, ns, nf, True)
    y, yd = ref.render(p, x, want_f64=True)
    bp = gpu.BatchProcessor("EasyExpander", ns, precision="f32").prepare(p[0])
    out = bp.process(x, params=p)
    r = out.astype(np.float64) - yd
    per_stream = np.array([dbfs_rms(r[s]) for s in range(ns)])
    print("fp32 residual dBFS rms per stream: min %.1f median %.1f max %.1f" % (per_stream.min(), np.median(per_stream), per_stream.max()))
    assert per_stream.max() < -100.0, per_stream.max()
    bp.close()


def test_matches_golden_from_reference(gpu):
    g = np.load(GOLD / "EasyExpander_auto.npz")
    ns, nf, block = int(g["n_streams"]), int(g["n_frames"]), int(g["block"])
    x = wl.programme_bursts(ns, nf)
    p = g["params"]
    bp = gpu.BatchProcessor("EasyExpander", ns, block_len=block, precision="f64").prepare(p[0])
    tap = np.zeros((ns, 2, nf))
    bp.process(x, tap64=tap, params=p)
    n = g["y64_head"].shape[-1]
    for got, want in ((tap[:, :, :n], g["y64_head"]), (tap[:, :, -n:], g["y64_tail"])):
        assert (np.abs(got - want) <= TOL64 * np.maximum(np.abs(want), np.abs(want).max(axis=2, keepdims=True))).all()
    assert np.allclose(np.abs(tap).sum(axis=2), g["y64_abs_sum"], rtol=1e-11, atol=0)
    bp.close()


def test_device_buffers_large_batch(gpu, ref):
    """4096 stereo streams x 1 s from device buffers (cfg3 shape), 10 streams re-rendered by the reference."""
    torch = pytest.importorskip("torch")
    ns, nf = 4096, 48000
    gsel = np.random.default_rng(3).choice(ns, 10, replace=False)
    xs = wl.programme_bursts(10, nf, first_stream=500)
    x = torch.zeros((ns, 2, nf), dtype=torch.float32, device="cuda")
    x[:] = torch.from_numpy(xs[0]).cuda()
    for i, s in enumerate(gsel):
        x[int(s)] = torch.from_numpy(xs[i]).cuda()
    nb = (nf + 511) // 512
    p = wl.automation(ref.sliders, nb, ns, which=[0, 1])
    bp = gpu.BatchProcessor("EasyExpander", ns, precision="f64").prepare(p[0])
    torch.cuda.synchronize()
    y = bp.process(x, params=p)
    bp.sync()
    yr, _ = ref.render(np.ascontiguousarray(p[:, gsel]), xs)
    got = y[torch.from_numpy(gsel).cuda()].cpu().numpy()
    assert np.mean(got == yr) > 0.9999
    assert np.abs(got.astype(np.float64) - yr).max() <= 2.0 ** -22 * np.abs(yr).max()
    bp.close()


def test_fp32_split_kernel_at_bench_shape_sampled(gpu, ref):
    """The fp32 split kernel at the cfg3 launch shape (4096 stereo streams -> 28 streams per CTA, device buffers, per-block
    automation through the device parameter path), 2 s; 24 streams spread over the batch re-rendered by the reference:
    residual of every sampled stream below -100 dBFS."""
    torch = pytest.importorskip("torch")
    ns, nf = 4096, 96000
    x = torch.empty((ns, 2, nf), dtype=torch.float32, device="cuda")
    gpu.synth_into(x, 1, 5000, 0)
    nb = (nf + 511) // 512
    k = 64
    a = wl.automation(ref.sliders, nb, k, which=[0, 1])
    p = np.ascontiguousarray(np.tile(a, (1, ns // k, 1)))
    bp = gpu.BatchProcessor("EasyExpander", ns, precision="f32").prepare(p[0])
    y = bp.process(x, params=p)
    bp.sync()
    pick = sorted(set(np.random.default_rng(11).choice(ns, 22, replace=False).tolist()) | {0, ns - 1})
    xs = wl.synth(1, 5000, pick, 2, nf)
    _, yd = ref.render(np.ascontiguousarray(p[:, pick]), xs, want_f64=True)
    got = y[torch.tensor(pick).cuda()].cpu().numpy().astype(np.float64)
    worst = max(dbfs_rms(got[i] - yd[i]) for i in range(len(pick)))
    print("fp32 split kernel, 4096 streams: worst sampled stream residual %.1f dBFS" % worst)
    assert worst < -100.0, worst
    assert bp.last_kernel_ms()[1] >= 1
    bp.close()


@pytest.mark.parametrize("filters,ns,nf,cut", [(False, 37, 48000 + 13, 20480), (True, 37, 30000 + 7, 512 * 9), (True, 5, 31, 0)])
def test_fp32_role_pipeline_v3_cases(gpu, ref, filters, ns, nf, cut):
    """The default fp32 kernel (k_easyexpander_v3.cu: block_len a multiple of 32): partial CTAs, a ragged last batch, the
    detector filters on and off, the state handed over between two calls (a checkpoint in between); against the float64
    reference (< -100 dBFS per stream) and against the round-1 split kernel (option scalar_f32 = 2)."""
    p, x = make_case(ref, ns, nf, True, detector_filters=filters)
    _, yd = ref.render(p, x, want_f64=True)
    outs = {}
    for variant in (0, 2):
        bp = gpu.BatchProcessor("EasyExpander", ns, precision="f32").prepare(p[0])
        bp.set_option("scalar_f32", variant)
        if cut:
            y0 = bp.process(np.ascontiguousarray(x[:, :, :cut]), params=np.ascontiguousarray(p[:cut // 512]))
            blob = bp.save_state()
            bp.close()
            bp = gpu.BatchProcessor("EasyExpander", ns, precision="f32").prepare(p[0])
            bp.set_option("scalar_f32", variant)
            bp.load_state(blob)
            y1 = bp.process(np.ascontiguousarray(x[:, :, cut:]), params=np.ascontiguousarray(p[cut // 512:]))
            outs[variant] = np.concatenate([y0, y1], axis=2)
        else:
            outs[variant] = bp.process(x, params=p)
        flag = bp.read_state("expanding")
        assert set(np.unique(flag)) <= {0.0, 1.0}
        for name in ("det_slow", "gain_db_s", "det_db_m", "gr_db_m", "hpf_lp", "lpf_y"):
            assert np.isfinite(bp.read_state(name)).all(), name
        bp.close()
    worst = max(dbfs_rms(outs[0][s].astype(np.float64) - yd[s]) for s in range(ns))
    between = max(dbfs_rms(outs[0][s].astype(np.float64) - outs[2][s]) for s in range(ns))
    print("fp32 v3: worst stream vs float64 reference %.1f dBFS; vs the split kernel %.1f dBFS" % (worst, between))
    assert worst < -100.0 and between < -100.0, (worst, between)


def test_fp32_v3_chunked_call_is_bit_identical(gpu, ref):
    """Scaled band states round-trip exactly through the float64 state rows: one call == the same frames in three calls."""
    ns, nf = 12, 512 * 30
    p, x = make_case(ref, ns, nf, True)
    bp = gpu.BatchProcessor("EasyExpander", ns, precision="f32").prepare(p[0])
    whole = bp.process(x, params=p)
    bp.close()
    bp = gpu.BatchProcessor("EasyExpander", ns, precision="f32").prepare(p[0])
    parts = []
    for b0, b1 in ((0, 7), (7, 8), (8, 30)):
        parts.append(bp.process(np.ascontiguousarray(x[:, :, b0 * 512:b1 * 512]), params=np.ascontiguousarray(p[b0:b1])))
    bp.close()
    assert np.array_equal(np.concatenate(parts, axis=2), whole)


def test_detector_filter_sliders_move_inside_a_call_on_the_device_path(gpu, ref):
    """All five sliders automated per block, the detector HPF / LPF frequencies included (the only @slider inputs that go
    through exp(): their coefficients come from per-grid-value tables made with the host libm).  The call must stay on the
    device parameter path (two launches: parameter kernel + render) and match the reference like any other render:
    fp64 <= 1e-12 of level, the hysteresis flag exact; fp32 below -100 dBFS."""
    ns, nf = 19, 512 * 60 + 77
    x = wl.programme_bursts(ns, nf)
    nb = (nf + 511) // 512
    base = np.array([s["default"] for s in ref.sliders], np.float32)
    base[3], base[4] = 120.0, 6000.0
    p = wl.automation(ref.sliders, nb, ns, which=[0, 1, 2, 3, 4], base=base)
    p[5:9, :, 3] = 0.0                                   # HPF off for a few blocks (hpf_on toggles), off-grid values elsewhere
    p[11, :, 4] = 20000.0                                # LPF off for one block
    assert len(np.unique(p[:, 0, 3])) > 10 and len(np.unique(p[:, 0, 4])) > 10
    y, yd, v = ref.render(p, x, want_f64=True, want_vars=True)
    bp = gpu.BatchProcessor("EasyExpander", ns, precision="f64").prepare(p[0])
    tap = np.zeros_like(yd)
    l0 = bp.launch_count()
    bp.process(x, tap64=tap, params=p)
    assert bp.launch_count() - l0 == 2                   # not the host coefficient path (one launch), not one launch per block
    assert rel_err(tap, yd).max() <= TOL64
    assert np.array_equal(bp.read_state("expanding"), v["expanding"])
    bp.close()
    bp = gpu.BatchProcessor("EasyExpander", ns, precision="f32").prepare(p[0])
    out = bp.process(x, params=p)
    assert max(dbfs_rms(out[s].astype(np.float64) - yd[s]) for s in range(ns)) < -100.0
    bp.close()


def test_long_automated_call_is_rendered_in_pipelined_segments(gpu, ref):
    """Calls of 512 host blocks or more are cut into time segments (parameter rows of segment k + 1 upload while segment k
    renders; state carries through the state rows): 600 blocks -> 2 segments = 4 launches, the result equal to the
    reference like any render (fp64 <= 1e-12 of level, flag exact) and, in float, bit-identical to the same frames fed
    in two calls at the segment boundary."""
    ns, nb = 6, 600
    nf = 512 * nb - 100                                  # ragged last block
    p, x = make_case(ref, ns, nf, True, detector_filters=True)
    y, yd, v = ref.render(p, x, want_f64=True, want_vars=True)
    bp = gpu.BatchProcessor("EasyExpander", ns, precision="f64").prepare(p[0])
    tap = np.zeros_like(yd)
    l0 = bp.launch_count()
    bp.process(x, tap64=tap, params=p)
    assert bp.launch_count() - l0 == 4
    assert rel_err(tap, yd).max() <= TOL64
    assert np.array_equal(bp.read_state("expanding"), v["expanding"])
    bp.close()
    bp = gpu.BatchProcessor("EasyExpander", ns, precision="f32").prepare(p[0])
    whole = bp.process(x, params=p)
    bp.close()
    bp = gpu.BatchProcessor("EasyExpander", ns, precision="f32").prepare(p[0])
    cut = 300 * 512
    parts = [bp.process(np.ascontiguousarray(x[:, :, :cut]), params=np.ascontiguousarray(p[:300])),
             bp.process(np.ascontiguousarray(x[:, :, cut:]), params=np.ascontiguousarray(p[300:]))]
    bp.close()
    assert np.array_equal(np.concatenate(parts, axis=2), whole)
    assert max(dbfs_rms(whole[s].astype(np.float64) - yd[s]) for s in range(ns)) < -100.0
