"""GPU parity tests for the Faust leaves RED, ClickBeGoneSG, VAR (generic stream-per-thread kernel, p_faust.cu) and GTS
(time-parallel kernel, k_gts.cu): the CUDA path through the C ABI against the float32 CPU restatements
(oracle/za_faust_port.c -- PARITY UNPINNED) on the same seeded inputs.
Criterion: fp32 residual below -100 dBFS RMS; parameter changes on the exact host block for any block length; state
carried across calls; ZA_F64 refused."""
import numpy as np
import pytest

import za_oracle as zo
from za_gpu import workloads as wl
from parity import dbfs_rms

pytestmark = pytest.mark.gpu
KEYS = ["RED", "ClickBeGoneSG", "VAR", "GTS"]


@pytest.fixture(scope="module")
def gpu(built_lib):
    import za_gpu
    if za_gpu.lib().za_gpu_device_count() < 1:
        pytest.fail("GPU test run without a CUDA device")
    return za_gpu


def make_input(key, port, ns, nf):
    x = wl.programme_bursts(ns, nf, n_ch=port.n_in)
    if key == "RED":                                                        # loud dense wet signal, reference mostly quiet: the leaf ducks
        x[:, 0:2] = 0.3 * np.sign(x[:, 0:2]) * np.abs(x[:, 0:2]) ** 0.3
        x[:, 4:6] *= np.where((np.arange(nf) // 12000) % 2 == 0, 0.001, 1.0).astype(np.float32)
    if key == "ClickBeGoneSG":                                              # needle clicks on programme material
        rng = np.random.default_rng(3)
        for s in range(ns):
            for f in (rng.choice(nf - 100, size=max(3, nf // 4000), replace=False) + 50 if nf > 200 else []):
                x[s, :, f] += rng.choice([-0.6, 0.6])
    return x


@pytest.mark.parametrize("key", KEYS)
@pytest.mark.parametrize("ns,nf,block,automate", [(37, 24000 + 77, 512, True), (5, 9000 + 1, 100, True), (70, 6000, 7, False), (1, 31, 512, False)])
def test_fp32_parity(gpu, key, ns, nf, block, automate):
    port = zo.PortPlugin(key)
    x = make_input(key, port, ns, nf)
    base = np.array([s["default"] for s in port.sliders], np.float32)
    if automate:
        which = [0, 1, 2] if key != "ClickBeGoneSG" else [0, 1, 2, 3]
        p = wl.automation(port.sliders, (nf + block - 1) // block, ns, which=which, base=base)
    else:
        p = np.tile(base, (ns, 1))
    y, _ = port.render(p, x, block=block)
    info = gpu.plugin_info(key)
    assert (info.n_in, info.n_out, info.kind) == (port.n_in, port.n_out, 1)
    bp = gpu.BatchProcessor(key, ns, 48000.0, block, "f32").prepare(p if p.ndim == 2 else p[0])
    out = bp.process(x, params=p if p.ndim == 3 else None)
    assert bp.launch_count() > 0
    r = out.astype(np.float64) - y
    print(f"{key} ns={ns} nf={nf} block={block}: residual {dbfs_rms(r):.1f} dBFS, equal {np.mean(out == y):.3f}")
    assert dbfs_rms(r) < -100.0, dbfs_rms(r)
    if nf >= 6000 and key != "VAR":
        assert dbfs_rms(out.astype(np.float64) - x[:, :port.n_out]) > -80.0   # the effect did something: not a vacuous comparison
    bp.close()


@pytest.mark.parametrize("key", KEYS)
def test_state_carries_across_calls_and_in_place(gpu, key):
    port = zo.PortPlugin(key)
    ns, nf, block = 9, 512 * 11 + 300, 512
    x = make_input(key, port, ns, nf)
    p = wl.automation(port.sliders, (nf + block - 1) // block, ns, which=[0, 1])
    y, _ = port.render(p, x, block=block)
    bp = gpu.BatchProcessor(key, ns, 48000.0, block, "f32").prepare(p[0])
    cuts = [0, 512 * 3, 512 * 8, nf]
    buf = x.copy()
    parts = []
    for a, b in zip(cuts[:-1], cuts[1:]):
        seg = np.ascontiguousarray(buf[:, :, a:b])
        parts.append(bp.process(seg, out=seg if port.n_in == port.n_out else None, params=p[a // block:]))      # in place
    out = np.concatenate(parts, axis=2)
    assert dbfs_rms(out.astype(np.float64) - y) < -100.0
    bp.prepare(p[0])
    whole = bp.process(x, params=p)
    assert dbfs_rms(whole.astype(np.float64) - out) < -130.0
    bp.close()


def test_f64_refused(gpu):
    for key in KEYS:
        with pytest.raises(gpu.ZaError, match="float32"):
            gpu.BatchProcessor(key, 4, precision="f64")


@pytest.mark.parametrize("ns,block,cuts", [(37, 512, (0, 31, 31 + 45, 31 + 45 + 4096 + 13, 12000)), (70, 100, (0, 6000, 6001, 9000)),
                                           (3, 7, (0, 64, 2000))])
def test_clickbegonesg_tiled_kernel_is_bit_identical_to_the_generic_kernel(gpu, ns, block, cuts):
    """cbg_tiled_kernel (p_faust.cu: predictors lane = frame out of a shared-memory ring, recurrences lane = stream) calls
    the same Cbg::predict / Cbg::tail as the generic stream-per-thread kernel (option scalar_f32 = 1): outputs AND every
    state word (delay-line ring, its position, the five followers) must be the same bits -- over several calls of ragged
    lengths (the ring is handed over through the state), host blocks that do not divide the 32-frame tiles, partial CTAs,
    and the Mode slider (which picks the predictor pair) moving per block."""
    key = "ClickBeGoneSG"
    port = zo.PortPlugin(key)
    nf = cuts[-1]
    x = make_input(key, port, ns, nf)
    base = np.array([s["default"] for s in port.sliders], np.float32)
    p = wl.automation(port.sliders, (nf + block - 1) // block, ns, which=[0, 1, 2, 3], base=base)
    outs, states = [], []
    for generic in (0, 1):
        bp = gpu.BatchProcessor(key, ns, 48000.0, block, "f32").prepare(p[0])
        bp.set_option("scalar_f32", generic)
        parts = []
        for c0, c1 in zip(cuts[:-1], cuts[1:]):
            if c0 % block:                       # a call starts a host block: only cut on block boundaries for the parameter rows
                b0 = c0 // block
                parts.append(bp.process(np.ascontiguousarray(x[:, :, c0:c1]), params=np.ascontiguousarray(p[b0:b0 + 1])))
            else:
                parts.append(bp.process(np.ascontiguousarray(x[:, :, c0:c1]), params=np.ascontiguousarray(p[c0 // block:(c1 + block - 1) // block])))
        outs.append(np.concatenate(parts, axis=2))
        states.append(np.stack([bp.read_state(f"state[{k}]") for k in range(70)]))
        bp.close()
    assert np.array_equal(outs[0].view(np.uint32), outs[1].view(np.uint32))
    assert np.array_equal(states[0][:64].astype(np.float32).view(np.uint32), states[1][:64].astype(np.float32).view(np.uint32))
    assert np.array_equal(states[0][65:].astype(np.float32).view(np.uint32), states[1][65:].astype(np.float32).view(np.uint32))
    assert np.abs(outs[0] - x[:, :2]).max() > 1e-3          # clicks were repaired: not a pass-through comparison


@pytest.mark.parametrize("ns,block,cuts", [(37, 512, (0, 31, 31 + 45, 31 + 45 + 4096 + 13, 12000)), (70, 100, (0, 6000, 6001, 9000)),
                                           (3, 7, (0, 64, 2000))])
@pytest.mark.parametrize("key,n_state", [("RED", 8), ("VAR", 27)])
def test_red_tiled_kernel_is_bit_identical_to_the_generic_kernel(gpu, key, n_state, ns, block, cuts):
    """red_tiled_kernel (p_faust.cu: powers, detector and gain lane = frame, the two + six followers lane = stream, five tiles
    in flight) is made of the same Red::powers / envelopes / detect / reduction_db / gain as Red::frame in the generic
    stream-per-thread kernel (option scalar_f32 = 1): six output channels AND the eight state words must be the same bits,
    over several calls of ragged lengths, host blocks that cut the 32-frame tiles, partial CTAs, all three sliders moving per
    block; the render is also checked against the restatement.  The same for var_tiled_kernel (Var::detector / control /
    envelope / bands / mix, three recurrence warps, 16 streams per CTA) and VAR's 27 state words (noise generator included)."""
    port = zo.PortPlugin(key)
    nf = cuts[-1]
    x = make_input(key, port, ns, nf)
    base = np.array([s["default"] for s in port.sliders], np.float32)
    p = wl.automation(port.sliders, (nf + block - 1) // block, ns, which=[0, 1, 2], base=base)
    outs, states, blobs = [], [], []
    for generic in (0, 1):
        bp = gpu.BatchProcessor(key, ns, 48000.0, block, "f32").prepare(p[0])
        bp.set_option("scalar_f32", generic)
        parts = []
        for c0, c1 in zip(cuts[:-1], cuts[1:]):
            if c0 % block:
                b0 = c0 // block
                parts.append(bp.process(np.ascontiguousarray(x[:, :, c0:c1]), params=np.ascontiguousarray(p[b0:b0 + 1])))
            else:
                parts.append(bp.process(np.ascontiguousarray(x[:, :, c0:c1]), params=np.ascontiguousarray(p[c0 // block:(c1 + block - 1) // block])))
        outs.append(np.concatenate(parts, axis=2))
        states.append(np.stack([bp.read_state(f"state[{k}]") for k in range(n_state)]))
        blobs.append(bp.save_state())
        bp.close()
    assert np.array_equal(outs[0].view(np.uint32), outs[1].view(np.uint32))
    assert np.array_equal(states[0], states[1], equal_nan=True)
    assert np.array_equal(blobs[0], blobs[1])                # the checkpoint blob: every state word as stored, integer ones included
    assert np.abs(outs[0][:, :2] - x[:, :2]).max() > 1e-4    # the effect did something: not a pass-through comparison
    if key == "RED":
        assert np.array_equal(outs[0][:, 2:], x[:, 2:])      # the other four channels pass through
    if all(c % block == 0 for c in cuts[1:-1]):              # one continuous automation: the restatement renders the same call
        y, _ = port.render(p, x, block=block)
        assert dbfs_rms(outs[0].astype(np.float64) - y) < -100.0


@pytest.mark.parametrize("key", ["VAR", "RED"])
def test_generic_kernel_streams_per_warp_does_not_change_a_bit(gpu, key):
    """faust_leaf_kernel with 8 / 16 / 32 streams per warp (option lanes_per_warp): identical bits,
    partial last warps and CTAs included."""
    port = zo.PortPlugin(key)
    ns, nf, block = 45, 6000 + 13, 512
    x = make_input(key, port, ns, nf)
    base = np.array([s["default"] for s in port.sliders], np.float32)
    p = wl.automation(port.sliders, (nf + block - 1) // block, ns, which=[0, 1, 2], base=base)
    y, _ = port.render(p, x, block=block)
    outs = {}
    for lanes in (32, 16, 8):
        bp = gpu.BatchProcessor(key, ns, 48000.0, block, "f32").prepare(p[0])
        bp.set_option("lanes_per_warp", lanes)
        bp.set_option("scalar_f32", 1)                   # the generic kernel (RED has a tiled one by default)
        outs[lanes] = (bp.process(x, params=p), np.stack([bp.read_state(f"state[{k}]") for k in range(8)]))
        bp.close()
    assert dbfs_rms(outs[32][0].astype(np.float64) - y) < -100.0
    for lanes in (16, 8):
        assert np.array_equal(outs[lanes][0].view(np.uint32), outs[32][0].view(np.uint32)), lanes
        assert np.array_equal(outs[lanes][1], outs[32][1], equal_nan=True), lanes
