#!/usr/bin/env python3
"""tools/emu_jsfx.py [KEY ...] -- compile the generated translation units (csrc/gen/g_<key>.cu) as plain C++
(-DZA_JSFX_HOST_EMU) and compare one-stream renders with the reference's compiled DSP (oracle/_ref) on the CPU.
Development check of the translator's semantics before GPU time is spent; also used by tests/test_jsfxc.py."""
import ctypes as C
import subprocess
import sys
from pathlib import Path

import numpy as np

REPO = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(REPO / "zorakaudio-experimental-plugins_b200"), str(REPO / "oracle")]
GEN = REPO / "zorakaudio-experimental-plugins_b200" / "csrc" / "gen"
OUT = REPO / "tools" / "_emu"


class GenLaunch(C.Structure):
    _fields_ = [("in_", C.c_void_p), ("out", C.c_void_p), ("tap", C.c_void_p),
                ("row_stride", C.c_int64), ("n_frames", C.c_int64), ("mem_slots", C.c_int64),
                ("n_streams", C.c_int), ("block_len", C.c_int), ("n_rows", C.c_int), ("n_decl", C.c_int),
                ("q", C.c_void_p), ("decl_index0", C.c_void_p),
                ("state", C.c_void_p), ("mem", C.c_void_p), ("mt", C.c_void_p), ("err", C.c_void_p), ("srate", C.c_double),
                ("fft_scratch", C.c_int64), ("tw", C.c_void_p)]


def build(key: str) -> Path:
    OUT.mkdir(exist_ok=True)
    so, src = OUT / f"libemu_{key}.so", GEN / f"g_{key}.cu"
    rt = GEN.parent / "jsfx_rt.cuh"
    if not so.exists() or so.stat().st_mtime < max(src.stat().st_mtime, rt.stat().st_mtime):
        subprocess.run(["g++", "-O1", "-ffp-contract=off", "-DZA_JSFX_HOST_EMU", "-x", "c++", "-shared", "-fPIC",
                        "-o", str(so), str(src)], check=True)
    return so


class Emu:
    def __init__(self, key, f64=True):
        self.lib = C.CDLL(str(build(key)))
        self.lib.emu_var_name.restype = C.c_char_p
        self.f64 = f64
        self.n_vars, self.n_in, self.n_out, self.mem_slots, self.fft_scratch = (self.lib.emu_info(i) for i in range(5))
        self.names = [self.lib.emu_var_name(k).decode() for k in range(self.n_vars)] if self.n_vars > 0 else []

    def render(self, sliders, params, x, block=512, srate=48000.0, mem_slots=None):
        """sliders: oracle layout (list of dicts); params [n_sliders] or [nb][n_sliders]; x [n_in][nf] -> y f32, tap f64, vars."""
        import za_oracle as zo
        dt = np.float64 if self.f64 else np.float32
        nf = x.shape[1]
        p = np.atleast_2d(np.asarray(params, np.float32))
        q = np.ascontiguousarray(zo.quantise(sliders, p))
        idx = np.array([s["index0"] for s in sliders], np.int32)
        slots = self.mem_slots if mem_slots is None else mem_slots          # (tests shrink the heap to provoke the overflow path)
        state = np.zeros(self.n_vars + 65, dt); mem = np.zeros(slots + 1 + self.fft_scratch, dt)   # + the spare slot out-of-range accesses land in + FFT scratch
        kk = np.arange(16384) * (2 * np.pi / 32768.0)
        tw = np.ascontiguousarray(np.stack([np.cos(kk), -np.sin(kk)], axis=1))
        mt = np.zeros(625, np.uint32); err = np.zeros(1, np.int32)
        xin = np.ascontiguousarray(x, np.float32)
        y = np.zeros((max(self.n_out, 1), nf), np.float32); tap = np.zeros((max(self.n_out, 1), nf), np.float64)
        a = GenLaunch(xin.ctypes.data, y.ctypes.data, tap.ctypes.data, nf, nf, slots, 1, block, 1, len(sliders),
                      q.ctypes.data, idx.ctypes.data, state.ctypes.data, mem.ctypes.data, mt.ctypes.data, err.ctypes.data, srate,
                      self.fft_scratch, tw.ctypes.data)
        self.lib.emu_prepare(C.byref(a), int(self.f64))
        a.n_rows = q.shape[0]
        self.lib.emu_process(C.byref(a), int(self.f64))
        assert err[0] == 0, "mem[] overflow"
        return y[:self.n_out], tap[:self.n_out], dict(zip(self.names, state[:self.n_vars]))


def compare(key, nf=48000, automate=False, verbose=True):
    import za_oracle as zo
    from za_gpu import workloads as wl
    ref = zo.RefPlugin(key)
    e = Emu(key)
    assert (e.n_in, e.n_out) == (ref.n_in, ref.n_out), ((e.n_in, e.n_out), (ref.n_in, ref.n_out))
    x = wl.programme_bursts(1, nf, n_ch=ref.n_in)
    base = np.array([s["default"] for s in ref.sliders], np.float32)
    if automate:
        nb = (nf + 511) // 512
        p = wl.automation(ref.sliders, nb, 1, which=list(range(min(3, len(base)))), base=base)
    else:
        p = base[None, :]
    pr = p if p.ndim == 3 else np.tile(base, (1, 1))
    y, yd, v = ref.render(pr, x, want_f64=True, want_vars=True)
    ye, te, ve = e.render(ref.sliders, p[:, 0] if p.ndim == 3 else p, x[0])
    d = np.abs(te - yd[0])
    pk = max(np.abs(yd[0]).max(), 1e-30)
    bad = [(n, ve[n], v[n][0]) for n in v if n in ve and abs(ve[n] - v[n][0]) > 1e-9 * max(1.0, abs(v[n][0]))]
    if verbose:
        first = int(np.argmax(d.max(axis=0) > 1e-12 * pk)) if (d.max() > 1e-12 * pk) else -1
        print(f"{key}: max|d|/peak {d.max() / pk:.3e} (peak {pk:.3g}) first frame over 1e-12: {first}; vars off: {len(bad)} {bad[:6]}")
    return d.max() / pk, bad


if __name__ == "__main__":
    keys = [a for a in sys.argv[1:] if not a.startswith("-")] or [l.split("(")[1].rstrip(")") for l in (GEN / "gen_list.h").read_text().split("X(")[1:] if False] 
    if not keys:
        import re
        keys = re.findall(r"X\((\w+)\)", (GEN / "gen_list.h").read_text().split("ZA_GEN_PLUGINS(X)")[1])
    for k in keys:
        try:
            compare(k, automate="--auto" in sys.argv)
        except Exception as ex:
            print(f"{k}: FAILED {type(ex).__name__}: {ex}")
