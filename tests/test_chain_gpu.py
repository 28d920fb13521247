"""GPU tests of serial plugin chains rendered on the device (za_chain_*, csrc/za_chain.cu): the chain
against the reference's compiled plugins run one after the other on the CPU, the float32 buffer of one being the input of
the next as in a host's plugin chain; side-chain routing as in RTT's 8-channel contract."""
import numpy as np
import pytest

import za_oracle as zo
from za_gpu import workloads as wl
from parity import dbfs_rms

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu(built_lib):
    import za_gpu
    if za_gpu.lib().za_gpu_device_count() < 1:
        pytest.fail("GPU test run without a CUDA device")
    return za_gpu


def oracle_chain(keys, params, x, block):
    """CPU: plugin k renders plugin k-1's float output; channels a later plugin wants beyond that come from the chain input."""
    cur = None
    for k, key in enumerate(keys):
        orc = zo.PortPlugin(key) if key in zo.FAUST_KEYS else zo.RefPlugin(key)
        if cur is None:
            xin = x[:, :orc.n_in]
        else:
            xin = np.zeros((x.shape[0], orc.n_in, x.shape[2]), np.float32)
            m = min(cur.shape[1], orc.n_in)
            xin[:, :m] = cur[:, :m]
            if orc.n_in > m:
                xin[:, m:] = x[:, m:orc.n_in]
        cur, _ = orc.render(params[k], np.ascontiguousarray(xin), block=block)
    return cur


@pytest.mark.parametrize("keys,chunk", [(["RTT", "EasyExpander"], 4), (["EasyExpander", "ModTilt", "cascade8"], 3), (["cascade8", "RTT"], 1),
                                        (["ERBTilt", "VAR", "GTS"], 16)])
def test_chain_equals_the_plugins_run_one_after_the_other(gpu, keys, chunk):
    block = 512
    ns, nf = 11, block * 13 + 200
    ch = gpu.Chain(keys, ns, 48000.0, block, "f64")
    ch.set_chunk_blocks(chunk)
    x = wl.programme_bursts(ns, nf, n_ch=ch.n_in)
    x[:, 2:] *= 0.05                                       # quiet keys: RTT has something to do
    params = []
    for key in keys:
        sl = [dict(default=float(s.def_), min=float(s.min), max=float(s.max)) for s in gpu.slider_layout(key)]
        base = wl.default_params(key, ns)
        p = wl.automation(sl, (nf + block - 1) // block, ns, which=[0, 1], base=base[0])
        p[:, :, 2:] = base[None, :, 2:]
        params.append(p)
    want = oracle_chain(keys, params, x, block)
    ch.prepare([p[0] for p in params])
    got = ch.process(x, params=params)
    assert got.shape == want.shape
    db = dbfs_rms(got.astype(np.float64) - want)
    print(f"chain {' -> '.join(keys)}: residual {db:.1f} dBFS, float-equal {np.mean(got == want):.4f}")
    assert db < -100.0
    if all(k not in zo.FAUST_KEYS and k != "cascade8" for k in keys):
        assert np.mean(got == want) > 0.999              # fp64 stages: the float32 hand-over is bit-identical almost everywhere
    # device buffers, split in two calls: state carries in every stage
    import torch
    ch.prepare([p[0] for p in params])
    xd = torch.from_numpy(x).cuda()
    cut = block * 5
    a = ch.process(xd[:, :, :cut].contiguous(), params=params)
    b = ch.process(xd[:, :, cut:].contiguous(), params=[p[5:] for p in params])
    ch.sync()
    both = torch.cat([a, b], dim=2).cpu().numpy()
    assert dbfs_rms(both.astype(np.float64) - got) < -140.0
    ch.close()
