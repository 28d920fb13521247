"""GPU tests of the host-buffer path of the C ABI (za_capi.cu::process_host): the 3-slot H2D / kernels / D2H pipeline
over SEVERAL time chunks, per-chunk parameter-row offsets, ragged lengths and the tap64 form must give the bits of a
single-chunk render; plus the argument refusals of za_gpu_create / za_gpu_process and the device generator's CPU twin."""
import numpy as np
import pytest

import za_oracle as zo
from za_gpu import workloads as wl
from parity import TOL64, dbfs_rms, rel_err

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu(built_lib):
    import za_gpu
    if za_gpu.lib().za_gpu_device_count() < 1:
        pytest.fail("GPU test run without a CUDA device")
    return za_gpu


def _case(gpu, key, ns, nf, block, automate):
    info = gpu.plugin_info(key)
    x = wl.noise_two_sines(ns, nf) if key == "cascade8" else wl.programme_bursts(ns, nf, n_ch=info.n_in)
    base = wl.default_params(key, ns)
    if not automate:
        return base, None, x
    sl = [dict(default=float(s.def_), min=float(s.min), max=float(s.max)) for s in gpu.slider_layout(key)]
    which = list(range(min(2, len(sl))))
    p = wl.automation(sl, (nf + block - 1) // block, ns, which=which, base=base[0])
    p[:, :, len(which):] = base[None, :, len(which):]
    return base, p, x


@pytest.mark.parametrize("key,prec", [("cascade8", "f64"), ("cascade8", "f32"), ("EasyExpander", "f32"), ("EasyExpander", "f64"),
                                      ("stft_mask2048", "f64"), ("RTT", "f64"), ("ModTilt", "f32"), ("ERBTilt", "f64"), ("ERBTilt", "f32")])
def test_multi_chunk_host_render_equals_single_chunk(gpu, key, prec):
    block = 512
    ns, nf = 9, block * 13 + 131                     # ragged: the last chunk and the last block are short
    p0, pauto, x = _case(gpu, key, ns, nf, block, automate=True)
    info = gpu.plugin_info(key)
    outs = []
    for chunk in (0, 2 * block, 5 * block):          # 0 = the default (one chunk for this size)
        bp = gpu.BatchProcessor(key, ns, 48000.0, block, prec).prepare(p0 if pauto is None else pauto[0])
        if chunk:
            bp.set_option("host_chunk_frames", chunk)
        tap = np.zeros((ns, info.n_out, nf), np.float64) if prec == "f64" else None
        y = bp.process(x, params=pauto, tap64=tap)
        outs.append((y.copy(), None if tap is None else tap.copy()))
        bp.close()
    for y, tap in outs[1:]:
        assert np.array_equal(y, outs[0][0]), key
        if tap is not None:
            assert np.array_equal(tap, outs[0][1]), key
    # and the one-chunk render is the reference's (JSFX leaves: oracle/_ref; ModTilt: the restatement)
    orc = zo.PortPlugin(key) if key == "ModTilt" else zo.best_oracle(key)
    yr, yrd = orc.render(pauto if pauto is not None else p0, x, want_f64=True, block=block)
    if prec == "f64":
        assert rel_err(outs[0][1], yrd).max() <= TOL64
    else:
        ref = yrd if yrd is not None else yr.astype(np.float64)
        assert dbfs_rms(outs[0][0].astype(np.float64) - ref) < -100.0


def test_error_inside_a_multi_chunk_call_leaves_no_copy_in_flight(gpu):
    """A translated leaf that runs past its heap fails the host call; every enqueued copy has drained when it returns."""
    ns, nf = 5, 512 * 8
    p0, _, x = _case(gpu, "SpectralStabilizer", ns, nf, 512, automate=False)
    bp = gpu.BatchProcessor("SpectralStabilizer.translated", ns, precision="f32")
    bp.set_option("heap_slots", 1000)
    bp.set_option("host_chunk_frames", 1024)
    y = np.zeros((ns, 2, nf), np.float32)
    with pytest.raises(gpu.ZaError):
        bp.prepare(p0)
        bp.process(x, out=y)
    snap = y.copy()
    bp.sync() if False else None
    assert np.array_equal(snap, y)
    bp.close()


def test_f32_refused_where_it_misses_the_criterion(gpu):
    for key in ("DDT", "BedRock"):
        with pytest.raises(gpu.ZaError, match="ZA_F64"):
            gpu.BatchProcessor(key, 4, precision="f32")
        gpu.BatchProcessor(key, 4, precision="f64").close()
    with pytest.raises(gpu.ZaError, match="float32"):
        gpu.BatchProcessor("ModTilt", 4, precision="f64")


def test_overlapping_buffers_refused_when_channel_counts_differ(gpu):
    ns, nf = 4, 1024
    bp = gpu.BatchProcessor("RTT", ns, precision="f64").prepare(None)
    buf = np.zeros((ns, 8, nf), np.float32)
    import ctypes as C
    rc = gpu.lib().za_gpu_process(bp._h, None, 0, buf.ctypes.data, buf.ctypes.data, nf, nf, gpu.ZA_MEM_HOST)
    assert rc == -1 and b"overlap" in gpu.lib().za_gpu_last_error()
    bp.close()
    # same channel counts: in place is legal
    bp = gpu.BatchProcessor("EasyExpander", ns, precision="f64").prepare(None)
    buf = wl.programme_bursts(ns, nf)
    want = bp.process(buf.copy())
    bp.prepare(None)
    rc = gpu.lib().za_gpu_process(bp._h, None, 0, buf.ctypes.data, buf.ctypes.data, nf, nf, gpu.ZA_MEM_HOST)
    assert rc == 0 and np.array_equal(buf, want)
    bp.close()


def test_render_one_shot_picks_the_leafs_precision(gpu):
    ns, nf = 3, 2048
    x = wl.programme_bursts(ns, nf)
    for key in ("ModTilt", "EasyExpander"):
        p = wl.default_params(key, ns)
        y = np.zeros((ns, 2, nf), np.float32)
        rc = gpu.lib().za_gpu_render(key.encode(), p.ctypes.data, x.ctypes.data, y.ctypes.data, ns, nf)
        assert rc == 0, gpu.lib().za_gpu_last_error()
        assert np.isfinite(y).all() and np.abs(y).max() > 0


@pytest.mark.parametrize("key", ["Roomalizer", "SOMA"])
def test_second_prepare_keeps_the_heap_like_the_reference(gpu, key):
    """prepareToPlay twice on one instance: resetStateStructOnly keeps st.mem (JSFXJuceProcessor.cpp:8972-8980), so the
    second render starts from the delay lines the first one left.  Checked against ONE reused reference instance."""
    ref = zo.RefPlugin(key)
    nf = 512 * 12
    x = wl.programme_bursts(1, nf, n_ch=ref.n_in)
    p = np.array([s["default"] for s in ref.sliders], np.float32)
    inst = zo.RefInstance(ref)
    bp = gpu.BatchProcessor(key, 1, precision="f64")
    for round_ in range(2):
        inst.prepare(p)
        _, yd = inst.process(p, x[0])
        bp.prepare(p[None])
        tap = np.zeros((1, ref.n_out, nf))
        bp.process(x, tap64=tap)
        assert rel_err(tap, yd[None]).max() <= TOL64, (key, round_)
    inst.close()
    bp.close()


def test_device_generator_equals_its_numpy_twin(gpu):
    import torch
    for kind, nch in ((0, 1), (1, 2), (1, 8)):
        ns, nf, first, t0 = 7, 10000, 123, 48000 * 3 + 17
        x = torch.empty((ns, nch, nf), dtype=torch.float32, device="cuda")
        gpu.synth_into(x, kind, 4242, first, t0)
        torch.cuda.synchronize()
        want = wl.synth(kind, 4242, range(first, first + ns), nch, nf, t0)
        assert np.array_equal(x.cpu().numpy(), want), kind
        cs, en = gpu.checksum(x)
        cs2, en2 = wl.checksum(want)
        assert np.array_equal(cs, cs2)
        assert np.allclose(en, en2, rtol=1e-9)


def test_two_devices_render_their_shards(gpu):
    """Streams shard over GPUs with no collective: handles on device 0 and 1 render contiguous stream ranges
    of one batch, both checked against the reference.  Needs two GPUs (run with `gpurun --gpus 2`)."""
    if gpu.lib().za_gpu_device_count() < 2:
        pytest.skip("one GPU visible")
    from za_gpu import shard
    ns, nf = 48, 4096 + 33
    ref = zo.RefPlugin("EasyExpander")
    x = wl.programme_bursts(ns, nf)
    base = np.array([s["default"] for s in ref.sliders], np.float32)
    p = wl.automation(ref.sliders, (nf + 511) // 512, ns, which=[0, 1], base=base)
    _, yd = ref.render(p, x, want_f64=True)
    for dev in (0, 1):
        a, b = shard.shard_range(ns, dev, 2)
        bp = gpu.BatchProcessor("EasyExpander", b - a, precision="f64", device=dev).prepare(p[0, a:b])
        tap = np.zeros((b - a, 2, nf))
        bp.process(np.ascontiguousarray(x[a:b]), params=np.ascontiguousarray(p[:, a:b]), tap64=tap)
        assert rel_err(tap, yd[a:b]).max() <= TOL64, dev
        bp.close()


@pytest.mark.parametrize("key,prec", [("cascade8", "f32"), ("cascade8", "f64"), ("EasyExpander", "f64"), ("stft_mask2048", "f32"),
                                      ("SOMA", "f64"), ("RTT", "f32"), ("ModTilt", "f32"), ("ERBTilt", "f32"), ("ERBTilt", "f64"), ("SpectralStabilizer", "f64"),
                                      ("VAR", "f32"), ("GTS", "f32"), ("ClickBeGoneSG", "f32")])
def test_checkpoint_resume_is_bit_identical(gpu, key, prec):
    """za_gpu_state_save / za_gpu_state_load: a render interrupted after its first part and resumed in a FRESH handle from
    the saved blob continues bit-identically (outputs and, for translated leaves, the rand()-driven state included)."""
    block = 512
    ns, nf = 7, block * 20 + 300
    cut = block * 9
    p0, pauto, x = _case(gpu, key, ns, nf, block, automate=True)
    info = gpu.plugin_info(key)
    bp = gpu.BatchProcessor(key, ns, 48000.0, block, prec).prepare(pauto[0])
    y1 = bp.process(np.ascontiguousarray(x[:, :, :cut]), params=pauto)
    blob = bp.save_state()
    y2 = bp.process(np.ascontiguousarray(x[:, :, cut:]), params=pauto[cut // block:])
    bp.close()
    fresh = gpu.BatchProcessor(key, ns, 48000.0, block, prec).prepare(None)      # defaults: everything must come from the blob
    fresh.load_state(blob)
    y2b = fresh.process(np.ascontiguousarray(x[:, :, cut:]), params=pauto[cut // block:])
    assert np.array_equal(y2, y2b), key
    # a blob from another batch shape is refused
    other = gpu.BatchProcessor(key, ns + 1, 48000.0, block, prec).prepare(None)
    with pytest.raises(gpu.ZaError, match="different plugin"):
        other.load_state(blob)
    other.close(); fresh.close()
    assert np.isfinite(y1).all() and info.n_out == y2.shape[1]


@pytest.mark.parametrize("key,factor", [("EasyExpander", 2), ("cascade8", 4), ("RTT", 2), ("ERBTilt", 8)])
def test_oversampling_like_the_reference_host(gpu, key, factor):
    """za_gpu_create_os: per host block linear up-sampling inside the block (its last frame held), the plugin at srate x factor on
    blocks of block_len x factor frames, box average back (src/JSFXJuceProcessor.cpp:8704-8789).  Checked against the reference's
    compiled plugin driven the same way on the CPU."""
    block = 512
    ns, nf = 7, block * 9 + 77
    p0, pauto, x = _case(gpu, key, ns, nf, block, automate=True)
    orc = zo.best_oracle(key)
    # the host's up-sampler, block-local, in float like upsampleLinearBlock (:8704-8727)
    xu = np.empty((ns, x.shape[1], nf * factor), np.float32)
    nxt = np.concatenate([x[:, :, 1:], x[:, :, -1:]], axis=2)
    last = ((np.arange(nf) + 1) % block == 0) | (np.arange(nf) == nf - 1)
    nxt[:, :, last] = x[:, :, last]
    for sub in range(factor):
        t = np.float32(sub) / np.float32(factor)
        xu[:, :, sub::factor] = x + (nxt - x) * t
    yu, _ = orc.render(pauto, xu, srate=48000.0 * factor, block=block * factor)
    want = (yu.astype(np.float64).reshape(ns, yu.shape[1], nf, factor).sum(axis=3) / factor).astype(np.float32)
    bp = gpu.BatchProcessor(key, ns, 48000.0, block, "f64", oversampling=factor).prepare(pauto[0])
    got = bp.process(x, params=pauto)
    db = dbfs_rms(got.astype(np.float64) - want)
    print(f"oversampling x{factor} {key}: residual {db:.1f} dBFS, float-equal {np.mean(got == want):.4f}")
    assert db < -140.0
    assert np.mean(got == want) > 0.99
    bp.close()


def test_one_call_several_gpus(gpu):
    """za_batch_*: one batch sharded over the visible GPUs (1 here unless run with `gpurun --gpus N`; with one GPU two shards on the
    same device exercise the same sharding, threading and row arithmetic) against the reference."""
    n_dev = gpu.lib().za_gpu_device_count()
    devices = list(range(n_dev)) if n_dev > 1 else [0, 0, 0]
    ns, nf = 50, 512 * 7 + 100
    ref = zo.RefPlugin("EasyExpander")
    x = wl.programme_bursts(ns, nf)
    base = np.array([s["default"] for s in ref.sliders], np.float32)
    p = wl.automation(ref.sliders, (nf + 511) // 512, ns, which=[0, 1], base=base)
    y, _ = ref.render(p, x)
    mb = gpu.MultiGpuBatch("EasyExpander", ns, devices, precision="f64").prepare(p[0])
    got = mb.process(x, params=p)
    assert np.mean(got == y) > 0.9999
    mb.close()
