"""GPU parity tests for the Faust leaf ModTilt (plugins/Dynamics/ModTilt/src/ModTilt.dsp): the CUDA path through the
C ABI against the float32 CPU restatement (oracle/za_faust_port.c -- PARITY UNPINNED: the Faust compiler's output is
not in the reference tree) on the same seeded inputs.

Criteria: fp32 residual vs the restatement below -100 dBFS RMS (BASELINE.json north_star); host-block positions of
parameter changes exact for any block length; the five filter states within float rounding; state carried across
calls; in == out legal; ZA_F64 refused (Faust leaves compute in float)."""
import numpy as np
import pytest

import za_oracle as zo
from za_gpu import workloads as wl

pytestmark = pytest.mark.gpu


def dbfs_rms(r):
    return 20 * np.log10(np.sqrt(np.mean(np.asarray(r, np.float64) ** 2)) + 1e-300)


@pytest.fixture(scope="module")
def gpu(built_lib):
    import za_gpu
    if za_gpu.lib().za_gpu_device_count() < 1:
        pytest.fail("GPU test run without a CUDA device")
    return za_gpu


@pytest.fixture(scope="module")
def port():
    return zo.PortPlugin("ModTilt")


def make_case(port, ns, nf, automate, block=512):
    x = wl.programme_bursts(ns, nf)
    nb = (nf + block - 1) // block
    base = np.array([-4.5, 3.0, 0.8], np.float32)
    if automate:
        return wl.automation(port.sliders, nb, ns, which=[0, 1, 2], base=base), x
    return np.tile(base, (ns, 1)), x


@pytest.mark.parametrize("automate,ns,nf,block,lanes", [
    (False, 21, 48000 + 13, 512, 0), (True, 37, 48000 + 5, 512, 8), (True, 70, 20000 + 7, 100, 16),
    (True, 33, 9000 + 1, 7, 32), (False, 1, 31, 512, 0), (True, 5, 4096, 64, 0)])
def test_fp32_parity(gpu, port, automate, ns, nf, block, lanes):
    p, x = make_case(port, ns, nf, automate, block)
    y, _ = port.render(p, x, block=block)
    bp = gpu.BatchProcessor("ModTilt", ns, 48000.0, block, "f32")
    if lanes:
        bp.set_option("warps_per_stream", lanes)
    bp.prepare(p if p.ndim == 2 else p[0])
    out = bp.process(x, params=p if p.ndim == 3 else None)
    assert bp.launch_count() > 0
    r = out.astype(np.float64) - y
    assert dbfs_rms(r) < -100.0, dbfs_rms(r)
    assert np.abs(r).max() <= 2e-5 * max(np.abs(y).max(), 1e-3)
    # the effect is not a pass-through at these settings (the comparison above is not vacuous)
    if nf >= 4096:
        assert dbfs_rms(out.astype(np.float64) - x) > -70.0
    bp.close()


def test_parameter_change_positions_exact(gpu, port):
    """Mix switches 1 -> 0 -> 1 at host-block starts; with mix = 0 the output is x * trim, with mix = 1 it is x * g *
    trim: a change applied one frame early or late shows up as a gain step at the wrong frame."""
    ns, block, nb = 9, 48, 40
    nf = block * nb
    x = wl.programme_bursts(ns, nf)
    p = np.tile(np.array([-6.0, 2.5, 1.0], np.float32), (nb, ns, 1))
    p[10:25, :, 2] = 0.0
    y, _ = port.render(p, x, block=block)
    bp = gpu.BatchProcessor("ModTilt", ns, 48000.0, block, "f32").prepare(p[0])
    out = bp.process(x, params=p)
    r = out.astype(np.float64) - y
    assert np.abs(r).max() <= 2e-5 * np.abs(y).max()
    for b in (10, 25):                       # residual right at the two switch frames is as small as anywhere else
        assert np.abs(r[:, :, b * block - 2:b * block + 2]).max() <= 2e-5 * np.abs(y).max()
    bp.close()


def test_state_carry_inplace_and_states(gpu, port):
    ns, nf = 40, 30000
    p, x = make_case(port, ns, nf, False)
    y, _ = port.render(p, x)
    bp = gpu.BatchProcessor("ModTilt", ns, precision="f32")
    bp.set_option("scalar_f32", 1)                        # the stream-per-thread kernel for every frame
    bp.prepare(p)
    cuts = [0, 1, 33, 5000, 5017, 20000, nf]
    parts = [bp.process(np.ascontiguousarray(x[:, :, a:b])) for a, b in zip(cuts[:-1], cuts[1:])]
    out = np.concatenate(parts, axis=2)
    assert dbfs_rms(out.astype(np.float64) - y) < -100.0
    bp.close()
    # one call and chunked calls are the same arithmetic on that kernel: bit-identical
    bp = gpu.BatchProcessor("ModTilt", ns, precision="f32")
    bp.set_option("scalar_f32", 1)
    bp.prepare(p)
    buf = x.copy()
    whole = bp.process(buf, out=buf)         # in place
    assert np.array_equal(whole, out)
    bp.close()
    # the default path (scan kernel for the whole host blocks, in place) re-associates chunk start states only
    bp = gpu.BatchProcessor("ModTilt", ns, precision="f32").prepare(p)
    buf = x.copy()
    whole = bp.process(buf, out=buf)
    assert dbfs_rms(whole.astype(np.float64) - out) < -120.0
    assert dbfs_rms(whole.astype(np.float64) - y) < -100.0
    env2 = bp.read_state("env2")
    assert env2.shape == (ns,) and np.all(env2 >= 0) and np.any(env2 > 0)
    bp.close()


def test_host_coefficients_match_restatement(gpu, port):
    """mix = 0 isolates trim; tilt = 0 makes g_lo = g_hi = 1 -> m2 = m -> ratio 1 -> output = input exactly."""
    ns, nf = 4, 6000
    x = wl.programme_bursts(ns, nf)
    p = np.tile(np.array([0.0, 3.0, 1.0], np.float32), (ns, 1))
    bp = gpu.BatchProcessor("ModTilt", ns, precision="f32").prepare(p)
    out = bp.process(x)
    y, _ = port.render(p, x)
    assert np.abs(out.astype(np.float64) - y).max() <= 1e-6
    assert np.abs(out - x).max() <= 1e-5
    bp.close()


def test_f64_refused(gpu):
    with pytest.raises(Exception) as e:
        gpu.BatchProcessor("ModTilt", 4, precision="f64")
    assert "float32" in str(e.value)


def test_device_parameter_path_equals_host_path(gpu, port):
    """Automation rows go through the device quantiser + host-built tables by default; `force_generic_io` takes the
    host quantiser + host libm per row.  Same grid index, same table entry -> identical output bits."""
    ns, nf, block = 50, 24000 + 3, 64
    nb = (nf + block - 1) // block
    x = wl.programme_bursts(ns, nf)
    p = wl.automation(port.sliders, nb, ns, which=[0, 1, 2], base=np.array([-3.0, 3.0, 0.7], np.float32))
    rng = np.random.default_rng(5)
    p[::7] += rng.uniform(-8, 8, p[::7].shape).astype(np.float32)          # out-of-range and off-grid host values
    outs = []
    for force in (0, 1):
        bp = gpu.BatchProcessor("ModTilt", ns, 48000.0, block, "f32")
        bp.set_option("force_generic_io", force)
        bp.prepare(p[0])
        outs.append(bp.process(x, params=p))
        bp.close()
    assert np.array_equal(outs[0], outs[1])
    y, _ = port.render(p, x, block=block)
    assert dbfs_rms(outs[0].astype(np.float64) - y) < -100.0


@pytest.mark.parametrize("ns,nf,block", [(37, 512 * 40 + 77, 512), (5, 256 * 9, 256), (130, 1024 * 6 + 1023, 1024), (3, 511, 512)])
def test_scan_kernel_against_restatement_and_scalar_kernel(gpu, port, ns, nf, block):
    """The time-parallel scan kernel (k_modtilt_scan.cu: one warp per stream, 256-frame tiles) renders the whole host
    blocks of a call, the stream-per-thread kernel the ragged tail: against the restatement (criterion) and against the
    stream-per-thread kernel alone (option scalar_f32), with every parameter moving per block; states handed over."""
    p, x = make_case(port, ns, nf, True, block)
    y, _ = port.render(p, x, block=block)
    outs, states = {}, {}
    for scalar in (0, 1):
        bp = gpu.BatchProcessor("ModTilt", ns, 48000.0, block, "f32")
        bp.set_option("scalar_f32", scalar)
        bp.prepare(p[0])
        # two calls: the second starts from the state the first left (scan kernel -> tail kernel -> scan kernel)
        cut = (nf // 2) // block * block + (block // 2 if nf > 2 * block else 0)
        o = np.zeros_like(x)
        if cut > 0:
            o[:, :, :cut] = bp.process(np.ascontiguousarray(x[:, :, :cut]), params=p)
        if cut % block == 0:
            o[:, :, cut:] = bp.process(np.ascontiguousarray(x[:, :, cut:]), params=p[cut // block:])
        else:                                    # a call boundary inside a host block restarts the block grid: render in one go
            bp.prepare(p[0])
            o = bp.process(x, params=p)
        outs[scalar] = o
        states[scalar] = np.stack([bp.read_state(n) for n in ("env2", "base", "lp_piv", "ratio", "mean_rdb")])
        bp.close()
    for scalar in (0, 1):
        r = outs[scalar].astype(np.float64) - y
        assert dbfs_rms(r) < -100.0, (scalar, dbfs_rms(r))
    d = outs[0].astype(np.float64) - outs[1]
    print(f"scan vs scalar kernel: {dbfs_rms(d):.1f} dBFS, equal {np.mean(outs[0] == outs[1]):.3f}")
    assert dbfs_rms(d) < -120.0
    # states agree to the accuracy of the kernel's MUFU maps (mean_rdb is in dB: 1e-5 dB is a gain of 1e-6)
    assert np.allclose(states[0], states[1], rtol=5e-4, atol=1e-5)
