"""Seeded synthetic workloads of BASELINE.json configs.

numpy generators for parity-sized cases (tests, golden fixtures, CPU baseline samples) and the
matching description of the full-size cases bench.py generates on the device.  No reference or
oracle code is used here: these are inputs only.
"""
from __future__ import annotations

import numpy as np

SRATE = 48000.0


def cascade8_params(n_streams: int, first_stream: int = 0) -> np.ndarray:
    """cfg2: per-stream fc log-uniform 40 Hz-16 kHz, Q uniform 0.5-4, seed 1000+stream.
    Column order = slider order of cascade8.jsfx (fc1,q1,...,fc6,q6,fc7,fc8,mix)."""
    p = np.empty((n_streams, 15), np.float32)
    lo = np.array([40.0, 4000.0, 40.0, 40.0, 4000.0, 40.0, 8000.0, 8000.0])      # HP LP BP BP LP HP 1p 1p
    hi = np.array([400.0, 16000.0, 16000.0, 16000.0, 16000.0, 400.0, 16000.0, 16000.0])
    for i in range(n_streams):
        r = np.random.default_rng(1000 + first_stream + i)
        fc = np.exp(r.uniform(np.log(lo), np.log(hi)))   # log-uniform inside 40 Hz-16 kHz, EQ-shaped:
        q = r.uniform(0.5, 4.0, 6)                       # high-passes low, low-passes high, bands anywhere
        p[i, 0:12:2] = fc[:6]
        p[i, 1:12:2] = q
        p[i, 12], p[i, 13] = fc[6], fc[7]
        p[i, 14] = r.uniform(0.0, 1.0)
    return p


def noise_two_sines(n_streams: int, n_frames: int, first_stream: int = 0, n_ch: int = 1) -> np.ndarray:
    """cfg2 input: seeded uniform noise x0.25 + two sines (-18 dBFS each); [stream][ch][frame] f32."""
    x = np.empty((n_streams, n_ch, n_frames), np.float32)
    t = np.arange(n_frames, dtype=np.float64) / SRATE
    for i in range(n_streams):
        r = np.random.default_rng(5000 + first_stream + i)
        f1, f2 = r.uniform(60.0, 400.0), r.uniform(1000.0, 6000.0)
        for c in range(n_ch):
            nz = r.uniform(-0.25, 0.25, n_frames)
            x[i, c] = (nz + 0.125 * np.sin(2 * np.pi * f1 * t + c) + 0.125 * np.sin(2 * np.pi * f2 * t)).astype(np.float32)
    return x


def programme_bursts(n_streams: int, n_frames: int, first_stream: int = 0, n_ch: int = 2) -> np.ndarray:
    """cfg3 input: noise x slow random envelope sweeping -60..-6 dBFS so gain computers traverse
    knee / hysteresis regions; [stream][ch][frame] f32."""
    x = np.empty((n_streams, n_ch, n_frames), np.float32)
    seg = 2400  # 50 ms envelope knots
    nk = n_frames // seg + 2
    for i in range(n_streams):
        r = np.random.default_rng(9000 + first_stream + i)
        knots_db = r.uniform(-60.0, -6.0, nk)
        env_db = np.interp(np.arange(n_frames) / seg, np.arange(nk), knots_db)
        env = 10.0 ** (env_db / 20.0)
        mono = r.uniform(-1.0, 1.0, n_frames)
        for c in range(n_ch):
            side = r.uniform(-1.0, 1.0, n_frames) * 0.3
            x[i, c] = (env * (mono + side) * 0.7).astype(np.float32)
    return x


def automation(sliders, n_blocks: int, n_streams: int, which, first_stream: int = 0, base=None) -> np.ndarray:
    """cfg3 per-block automation: sliders listed in `which` (column indices) follow a seeded
    piecewise-linear + sine-LFO trajectory sampled at block starts; others stay at `base`
    (default: declared defaults).  -> float32 [n_blocks][n_streams][n_sliders] of HOST values."""
    ns = len(sliders)
    p = np.empty((n_blocks, n_streams, ns), np.float32)
    dflt = np.array([s["default"] for s in sliders], np.float32) if base is None else np.asarray(base, np.float32)
    p[:] = dflt
    b = np.arange(n_blocks, dtype=np.float64)
    for i in range(n_streams):
        r = np.random.default_rng(7000 + first_stream + i)
        for j in which:
            lo, hi = sliders[j]["min"], sliders[j]["max"]
            nk = max(2, n_blocks // 40 + 2)
            knots = r.uniform(lo + 0.1 * (hi - lo), hi - 0.1 * (hi - lo), nk)
            traj = np.interp(b / 40.0, np.arange(nk), knots)
            traj += 0.05 * (hi - lo) * np.sin(2 * np.pi * b / r.uniform(13.0, 90.0))
            p[:, i, j] = np.clip(traj, lo, hi).astype(np.float32)
    return p


def default_params(plugin_key: str, n_streams: int, first_stream: int = 0) -> np.ndarray:
    """Static per-stream parameter block of a plugin's bench workload: cascade8 -> the seeded cfg2
    parameters; catalog leaves -> declared slider defaults for every stream."""
    if plugin_key == "cascade8":
        return cascade8_params(n_streams, first_stream)
    import za_gpu
    d = np.array([s.def_ for s in za_gpu.slider_layout(plugin_key)], np.float32)
    return np.tile(d, (n_streams, 1))


def cfg1_input(n_frames: int = 480000, n_ch: int = 2, srate: float = 48000.0, seed: int = 12345) -> np.ndarray:
    """BASELINE config 1 input: seeded LCG white noise in [-0.5, 0.5) gated 100 ms on / 200 ms
    off plus a 1 kHz sine at -12 dBFS; one stream, [1][ch][frame] f32 (channel c uses seed + c)."""
    x = np.empty((1, n_ch, n_frames), np.float32)
    t = np.arange(n_frames)
    gate = (t % int(0.3 * srate)) < int(0.1 * srate)
    sine = 10.0 ** (-12.0 / 20.0) * np.sin(2.0 * np.pi * 1000.0 * t / srate)
    for c in range(n_ch):
        st = np.uint64(seed + c)
        a, inc, mask = np.uint64(6364136223846793005), np.uint64(1442695040888963407), np.uint64(0xFFFFFFFFFFFFFFFF)
        # 64-bit LCG advanced in closed form per frame: state_n = a^n s + inc (a^n - 1)/(a - 1); done iteratively in
        # chunks with Python ints would be slow, so jump tables of (a^k, c_k) for k = 1 .. chunk are built once
        chunk = 4096
        ak, ck = np.empty(chunk, np.uint64), np.empty(chunk, np.uint64)
        pa, pc = 1, 0
        for k in range(chunk):
            pa, pc = (pa * int(a)) & int(mask), (pc * int(a) + int(inc)) & int(mask)
            ak[k], ck[k] = pa, pc
        noise = np.empty(n_frames, np.float64)
        cur = int(st)
        with np.errstate(over="ignore"):
            for f0 in range(0, n_frames, chunk):
                n = min(chunk, n_frames - f0)
                stt = ak[:n] * np.uint64(cur) + ck[:n]
                noise[f0:f0 + n] = (stt >> np.uint64(40)).astype(np.float64) / float(1 << 24) - 0.5
                cur = int(stt[n - 1])
        x[0, c] = (noise * gate + sine).astype(np.float32)
    return x


# ---- the library's device generator (csrc/k_synth.cu), restated in numpy: bit-identical by construction ------------------
def _mix32(x):
    x = np.asarray(x, np.uint32).copy()
    x ^= x >> np.uint32(16); x *= np.uint32(0x7feb352d); x ^= x >> np.uint32(15); x *= np.uint32(0x846ca68b); x ^= x >> np.uint32(16)
    return x


def _stream_key(seed, stream, ch):
    with np.errstate(over="ignore"):
        return _mix32(np.uint32(seed) * np.uint32(0x9E3779B1) + np.asarray(stream, np.uint32) * np.uint32(0x85EBCA77)
                      + np.uint32(ch) * np.uint32(0xC2B2AE3D) + np.uint32(0x27D4EB2F))


def _uni(h):
    return (h >> np.uint32(8)).astype(np.float32) * np.float32(2.0 ** -23) + np.float32(-1.0)


def _frame_hash(key, t):
    lo = (t & 0xFFFFFFFF).astype(np.uint32)
    hi = (t >> 32).astype(np.uint32)
    with np.errstate(over="ignore"):
        return _mix32((lo * np.uint32(0x9E3779B1) + hi * np.uint32(0x632BE5AB)) ^ key)


_TABLES = None


def synth_tables():
    global _TABLES
    if _TABLES is None:
        import za_gpu
        sine, amp = np.empty(4096, np.float32), np.empty(256, np.float32)
        za_gpu._check(za_gpu.lib().za_gpu_synth_tables(sine.ctypes.data, amp.ctypes.data), "za_gpu_synth_tables")
        _TABLES = (sine, amp)
    return _TABLES


def synth(kind: int, seed: int, streams, n_ch: int, n_frames: int, t0: int = 0) -> np.ndarray:
    """CPU twin of za_gpu_synth for the listed absolute stream numbers -> float32 [len(streams)][n_ch][n_frames]."""
    sine, amp = synth_tables()
    streams = list(streams)
    out = np.empty((len(streams), n_ch, n_frames), np.float32)
    t = np.arange(t0, t0 + n_frames, dtype=np.int64)
    t32 = (t & 0xFFFFFFFF).astype(np.uint32)
    with np.errstate(over="ignore"):
        for a, s in enumerate(streams):
            key_m = _stream_key(seed, s, 255)
            if kind == 0:
                inc1 = np.uint32(5368709 + int(_mix32(key_m ^ np.uint32(0x1111))) % 30422000)
                inc2 = np.uint32(89478485 + int(_mix32(key_m ^ np.uint32(0x2222))) % 447392000)
                s1 = sine[(t32 * inc1) >> np.uint32(20)]
                s2 = sine[(t32 * inc2) >> np.uint32(20)]
            else:
                k = t // 2400
                r = (t - k * 2400).astype(np.float32)
                k32 = (k & 0xFFFFFFFF).astype(np.uint32)
                a0 = amp[_mix32(key_m ^ np.uint32(0xA5A5A5A5) ^ (k32 * np.uint32(0x9E3779B1))) & np.uint32(255)]
                a1 = amp[_mix32(key_m ^ np.uint32(0xA5A5A5A5) ^ ((k32 + np.uint32(1)) * np.uint32(0x9E3779B1))) & np.uint32(255)]
                e = a0 + (a1 - a0) * (r * np.float32(4.1666667675599456e-04))
                mono = _uni(_frame_hash(key_m, t))
            for c in range(n_ch):
                key = _stream_key(seed, s, c)
                u = _uni(_frame_hash(key, t))
                if kind == 0:
                    out[a, c] = (np.float32(0.25) * u + np.float32(0.125) * s1) + np.float32(0.125) * s2
                else:
                    out[a, c] = e * (mono + np.float32(0.3) * u)
    return out


def checksum(x: np.ndarray):
    """CPU twin of za_gpu_checksum: per-row wrapping-uint64 checksum and energy of float32 [rows...][frames]."""
    x = np.ascontiguousarray(x, np.float32)
    nf = x.shape[-1]
    with np.errstate(over="ignore"):
        w = (_mix32(np.arange(nf, dtype=np.uint32)) | np.uint32(1)).astype(np.uint64)
        bits = x.view(np.uint32).astype(np.uint64) + np.uint64(1)
        s = (bits * w).sum(axis=-1, dtype=np.uint64)
    return s, (x.astype(np.float64) ** 2).sum(axis=-1)
