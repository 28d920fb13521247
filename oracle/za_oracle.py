"""oracle/za_oracle.py -- TEST INFRASTRUCTURE ONLY.

Python harness around (a) oracle/_ref/libref_<key>.so = the reference's own compiled DSP-JSFX
object behind oracle/ref_shim.c, and (b) oracle/libza_oracle_port.so = the plain-C restatement
(oracle/za_oracle_port.c).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this module; the product (libza_gpu.so) never does.

Also holds the seeded synthetic signal / parameter generators of SURVEY.md section 8(d) so the
tests, the golden-fixture generator and bench.py all render exactly the same inputs.
"""
from __future__ import annotations

import ctypes as C
import json
import os
import re
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
REPO = HERE.parent
PKG = REPO / "zorakaudio-experimental-plugins_b200"
REF_DIR = HERE / "_ref"


# ------------------------------------------------------------------------------------------------
# slider layout (restates parseJsfxSliderDecls, src/JSFXJuceProcessor.cpp:710-939, numeric part)
# ------------------------------------------------------------------------------------------------
class SliderLayout(C.Structure):
    _fields_ = [("n", C.c_int32), ("index0", C.c_int32 * 64), ("alias", C.c_int32 * 64),
                ("vmin", C.c_float * 64), ("vmax", C.c_float * 64), ("vstep", C.c_float * 64),
                ("is_choice", C.c_int32 * 64)]


_RE_SLIDER = re.compile(r"^\s*slider\s*([0-9]{1,2})\s*:\s*([^<\r\n;]+)\s*(?:<\s*([^>]+)\s*>)?\s*(.*)$")


def _parse_float(tok: str):
    m = re.match(r"\s*[-+]?(?:\d+\.?\d*(?:[eE][-+]?\d+)?|\.\d+(?:[eE][-+]?\d+)?)", tok)
    if not m:
        return None
    return np.float32(float(m.group(0)))


def _split_top_commas(s: str):
    parts, cur, depth = [], "", 0
    for ch in s:
        if ch == "{":
            depth += 1
        elif ch == "}" and depth > 0:
            depth -= 1
        if ch == "," and depth == 0:
            parts.append(cur.strip()); cur = ""
        else:
            cur += ch
    parts.append(cur.strip())
    return parts


def parse_sliders(jsfx_text: str):
    """-> list of dict(index0, var, default, min, max, step, is_choice), sorted by index."""
    out = {}
    for line in re.split(r"[\r\n]+", jsfx_text):
        m = _RE_SLIDER.match(line)
        if not m:
            continue
        n = int(m.group(1))
        if n < 1 or n > 64:
            continue
        tok = m.group(2).strip()
        var, deftok = "", tok
        if "=" in tok:
            eq = tok.rfind("=")
            var, deftok = tok[:eq].strip(), tok[eq + 1:].strip()
        d = _parse_float(deftok)
        d = np.float32(0.0) if d is None else d
        rng = m.group(3)
        if rng is not None and rng.strip().lower() in ("string", "str", "text"):
            continue
        if var.startswith("#"):
            continue
        vmin, vmax, vstep, is_choice = np.float32(0), np.float32(1), np.float32(0.001), False
        if rng is not None:
            parts = _split_top_commas(rng)
            if len(parts) >= 2:
                a, b = _parse_float(parts[0]), _parse_float(parts[1])
                vmin = np.float32(0) if a is None else a
                vmax = np.float32(1) if b is None else b
            if len(parts) >= 3:
                st = parts[2]
                if "{" in st:
                    bp = st.find("{"); cp = st.find("}", bp + 1)
                    if cp >= 0 and [x for x in st[bp + 1:cp].split(",") if x.strip()]:
                        is_choice = True
                    st = st[:bp].strip()
                if ":" in st:
                    st = st[:st.find(":")].strip()
                if st == "":
                    vstep = np.float32(1)
                else:
                    v = _parse_float(st)
                    vstep = np.float32(1) if v is None else v
            if vmax < vmin:
                vmin, vmax = vmax, vmin
            if not (vstep > 0):
                vstep = np.float32(0.001)
            d = min(max(d, vmin), vmax)
        if (n - 1) not in out:
            out[n - 1] = dict(index0=n - 1, var=var, default=float(d), min=float(vmin), max=float(vmax),
                              step=float(vstep), is_choice=is_choice)
    return [out[k] for k in sorted(out)]


def make_layout(sliders, var_index) -> SliderLayout:
    L = SliderLayout()
    L.n = len(sliders)
    for j, s in enumerate(sliders):
        L.index0[j] = s["index0"]
        L.alias[j] = var_index(s["var"]) if s["var"] else -1
        L.vmin[j], L.vmax[j], L.vstep[j] = s["min"], s["max"], s["step"]
        L.is_choice[j] = 1 if s["is_choice"] else 0
    return L


def quantise(sliders, values: np.ndarray) -> np.ndarray:
    """numpy restatement of hostParameterToJsfxSliderValue (src/JSFXJuceProcessor.cpp:5556-5597);
    values[..., n_sliders] float32 -> float64."""
    v = np.asarray(values, dtype=np.float32).astype(np.float64)
    out = np.empty_like(v)
    for j, s in enumerate(sliders):
        mn, mx, st = float(np.float32(s["min"])), float(np.float32(s["max"])), float(np.float32(s["step"]))
        x = v[..., j]
        if s["is_choice"]:
            x = mn + _llround(x) * st
        x = np.clip(x, mn, mx)
        if not s["is_choice"] and st > 0:
            x = np.clip(mn + _llround((x - mn) / st) * st, mn, mx)
        out[..., j] = x
    return out


def _llround(x):
    return np.where(x >= 0, np.floor(x + 0.5), -np.floor(-x + 0.5))


# ------------------------------------------------------------------------------------------------
# plugin sources (the catalog stays where it is; authored workloads live in the package)
# ------------------------------------------------------------------------------------------------
class _RefSliderDecl(C.Structure):
    _fields_ = [("index0", C.c_int32), ("is_choice", C.c_int32), ("n_choices", C.c_int32), ("is_string", C.c_int32), ("hidden", C.c_int32),
                ("shape", C.c_int32), ("default", C.c_float), ("min", C.c_float), ("max", C.c_float), ("step", C.c_float),
                ("shape_mod", C.c_float), ("var", C.c_char * 64), ("name", C.c_char * 128)]


def have_ref_parser() -> bool:
    return (HERE / "_ref" / "libref_parser.so").exists()


def ref_parse_sliders(text: str):
    """The REFERENCE'S OWN parseJsfxSliderDecls (src/JSFXJuceProcessor.cpp:710-939), compiled from where it lies by
    build_ref.build_parser() behind oracle/ref_parser_shim.cpp: list of dicts like parse_sliders() below."""
    lib = C.CDLL(str(HERE / "_ref" / "libref_parser.so"))
    lib.ref_parse_sliders.argtypes = [C.c_char_p, C.POINTER(_RefSliderDecl), C.c_int]
    arr = (_RefSliderDecl * 64)()
    n = lib.ref_parse_sliders(text.encode("utf-8", "replace"), arr, 64)
    return [dict(index0=d.index0, default=d.default, min=d.min, max=d.max, step=d.step, is_choice=bool(d.is_choice), n_choices=d.n_choices,
                 is_string=bool(d.is_string), hidden=bool(d.hidden), shape=d.shape, shape_mod=d.shape_mod, var=d.var.decode(),
                 name=d.name.decode("utf-8", "replace")) for d in arr[:n]]


def plugin_source_text(key: str) -> str:
    """Slider layout text.  Authored workloads: read the .jsfx.  Catalog leaves: the layout is
    committed in the package's plugin table (the reference tree does not exist on the GPU box)."""
    for p in (PKG / "plugins").rglob("*.jsfx"):
        if p.parent.parent.name == key:
            return p.read_text()
    if key in CATALOG_SLIDERS:
        return CATALOG_SLIDERS[key]
    cat = HERE / "catalog_sliders.json"        # written by tools/gen_plugins.py from the plugin sources
    if cat.exists():
        d = json.loads(cat.read_text())
        if key in d:
            return d[key]
    raise KeyError(key)


# slider declaration lines of the catalog leaves under test (their parameter API, restated from the
# plugin sources: EasyExpander.jsfx:13-21)
CATALOG_SLIDERS = {
    "EasyExpander": (
        "slider1:thresh_db=-40<-80,0,0.1>Threshold (dB)\n"
        "slider2:depth_db=24<0,60,1>Depth (dB)\n"
        "slider3:contour=50<0,100,1>Contour (0=gentle, 100=hard)\n"
        "slider4:det_hpf_hz=0<0,20000,1:log=1000>Detector HPF (Hz, 0=Off)\n"
        "slider5:det_lpf_hz=20000<20,20000,1:log=1000>Detector LPF (Hz, 20k=Off)\n"),
    # Faust leaf: hslider(label, init, min, max, step) in buildUserInterface order (ModTilt.dsp:9-19), written in the
    # slider-line form the library's parameter API uses
    "ModTilt": (
        "slider1:tilt_db=0<-6,3,0.1>Tilt (dB)\n"
        "slider2:pivot_hz=3<2,5,0.01>Pivot (Hz)\n"
        "slider3:mix=1<0,1,0.001>Mix\n"),
    "RED": (                                                   # RED.dsp:8-10
        "slider1:amount_db=12<0,24,0.1>Amount (max duck dB)\n"
        "slider2:sens_pct=50<0,100,1>Sensitivity (%)\n"
        "slider3:rel_ms=350<50,1200,1>Release (ms)\n"),
    "ClickBeGoneSG": (                                         # Click-Be-Gone (SG).dsp:11-15
        "slider1:amount=50<0,100,1>Amount [%]\n"
        "slider2:sensitivity=50<0,100,1>Sensitivity [%]\n"
        "slider3:hpf_hz=1500<300,6000,10>HPF [Hz]\n"
        "slider4:mode=1<0,2,1{Fast,Medium,Slow}>Mode\n"
        "slider5:monitor=0<0,1,1{Output,Delta}>Monitor\n"),
    "VAR": (                                                   # VAR.dsp:88-90
        "slider1:amount=35<0,100,1>Air Amount [%]\n"
        "slider2:sens=50<0,100,1>Sensitivity [%]\n"
        "slider3:floor_db=-60<-90,-30,1>Detector Floor [dB]\n"),
    "GTS": (                                                   # GTS.dsp:52-64
        "slider1:sigma_ms=2<0.1,8,0.01>Gaussian Sigma [ms]\n"
        "slider2:attack_db=0<-12,12,0.1>Attack Gain [dB]\n"
        "slider3:sustain_db=0<-12,12,0.1>Sustain Gain [dB]\n"
        "slider4:mix=1<0,1,0.01>Mix\n"
        "slider5:out_db=0<-12,12,0.1>Output Gain [dB]\n"),
}


class RefPlugin:
    """The reference's compiled DSP for one plugin (oracle/_ref/libref_<key>.so)."""

    def __init__(self, key: str, sliders=None):
        so = REF_DIR / f"libref_{key}.so"
        if not so.exists():
            raise FileNotFoundError(f"{so} (run oracle/build_ref.py where /root/reference exists)")
        self.key = key
        self.lib = C.CDLL(str(so))
        L = self.lib
        L.zref_var_index.argtypes = [C.c_char_p]
        L.zref_new.restype = C.c_void_p
        L.zref_new.argtypes = [C.POINTER(SliderLayout)]
        L.zref_free.argtypes = [C.c_void_p]
        L.zref_prepare.argtypes = [C.c_void_p, C.c_double, C.c_void_p]
        L.zref_vars.restype = C.POINTER(C.c_double)
        L.zref_vars.argtypes = [C.c_void_p]
        L.zref_mem.restype = C.POINTER(C.c_double)
        L.zref_mem.argtypes = [C.c_void_p, C.POINTER(C.c_int64)]
        L.zref_render.argtypes = [C.POINTER(SliderLayout), C.c_double, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                  C.c_void_p, C.c_int, C.c_int64, C.c_int64, C.c_int, C.c_int]
        L.zref_render_vars.argtypes = L.zref_render.argtypes + [C.c_void_p]
        self.n_in, self.n_out = L.zref_in_channels(), L.zref_out_channels()
        self.var_cap = L.zref_var_cap()
        self.sliders = sliders if sliders is not None else parse_sliders(plugin_source_text(key))
        self.layout = make_layout(self.sliders, lambda name: L.zref_var_index(name.encode()))
        self.meta = json.loads((REF_DIR / f"{key}.json").read_text()) if (REF_DIR / f"{key}.json").exists() else {}

    def render(self, params, x, srate=48000.0, block=512, want_f64=False, threads=0, want_vars=False):
        """params [n_streams,n_sliders] or [n_blocks,n_streams,n_sliders] f32; x [n_streams,n_in,n_frames] f32
        -> (y f32 [n_streams,n_out,n_frames], y64 or None[, {var name: final value per stream}])."""
        x = np.ascontiguousarray(x, dtype=np.float32)
        ns, nin, nf = x.shape
        assert nin == self.n_in, (nin, self.n_in)
        if want_vars:
            p = np.ascontiguousarray(params, dtype=np.float32)
            npb = 1 if p.ndim == 2 else p.shape[0]
            y = np.zeros((ns, self.n_out, nf), np.float32)
            yd = np.zeros((ns, self.n_out, nf), np.float64) if want_f64 else None
            vars_out = np.zeros((ns, self.var_cap), np.float64)
            rc = self.lib.zref_render_vars(C.byref(self.layout), srate, p.ctypes.data, npb, x.ctypes.data, y.ctypes.data,
                                           yd.ctypes.data if want_f64 else None, ns, nf, nf, block,
                                           threads or (os.cpu_count() or 1), vars_out.ctypes.data)
            assert rc == 0
            return y, yd, {k: vars_out[:, i] for k, i in self.meta.get("vars", {}).items()}
        p = np.ascontiguousarray(params, dtype=np.float32)
        npb = 1 if p.ndim == 2 else p.shape[0]
        assert p.shape[-2] == ns and p.shape[-1] == len(self.sliders)
        y = np.zeros((ns, self.n_out, nf), np.float32)
        yd = np.zeros((ns, self.n_out, nf), np.float64) if want_f64 else None
        th = threads or (os.cpu_count() or 1)
        rc = self.lib.zref_render(C.byref(self.layout), srate, p.ctypes.data, npb, x.ctypes.data, y.ctypes.data,
                                  yd.ctypes.data if want_f64 else None, ns, nf, nf, block, th)
        assert rc == 0
        return y, yd


class RefInstance:
    """ONE reference instance kept alive across prepare / process calls (zref_new / zref_prepare /
    zref_process_block_f64 of oracle/ref_shim.c): what a host does when it calls prepareToPlay twice on a plugin."""

    def __init__(self, ref: RefPlugin):
        self.ref, L = ref, ref.lib
        L.zref_process_block_f64.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int]
        self.h = L.zref_new(C.byref(ref.layout))

    def prepare(self, params, srate=48000.0):
        p = np.ascontiguousarray(params, np.float32)
        self.ref.lib.zref_prepare(self.h, srate, p.ctypes.data)

    def process(self, params, x, block=512):
        """params [n_sliders] or [n_blocks][n_sliders]; x [n_in][n_frames] -> (y f32, y f64) [n_out][n_frames]."""
        x = np.ascontiguousarray(x, np.float32)
        nch = max(self.ref.n_in, self.ref.n_out)
        nf = x.shape[1]
        p = np.ascontiguousarray(params, np.float32)
        p = p[None] if p.ndim == 1 else p
        y, yd = np.zeros((nch, nf), np.float32), np.zeros((nch, nf), np.float64)
        xin = np.zeros((nch, nf), np.float32)
        xin[:self.ref.n_in] = x
        PF, PD = C.POINTER(C.c_float), C.POINTER(C.c_double)
        for b, f0 in enumerate(range(0, nf, block)):
            n = min(block, nf - f0)
            ins = (PF * nch)(*[C.cast(xin[c, f0:].ctypes.data, PF) for c in range(nch)])
            outs = (PF * nch)(*[C.cast(y[c, f0:].ctypes.data, PF) for c in range(nch)])
            outd = (PD * nch)(*[C.cast(yd[c, f0:].ctypes.data, PD) for c in range(nch)])
            row = p[min(b, len(p) - 1)]
            self.ref.lib.zref_process_block_f64(self.h, row.ctypes.data, ins, outs, outd, nch, n)
        return y[:self.ref.n_out], yd[:self.ref.n_out]

    def close(self):
        if self.h:
            self.ref.lib.zref_free(self.h)
            self.h = None


class ShadowPlugin:
    """The SECONDARY oracle: the plugin's .jsfx source run by the reference's own WDL EEL2 interpreter (oracle/shadow_eel.cpp,
    built into oracle/_ref/libshadow_eel.so), driven like the reference's --correctness-check shadow VM.  One stream at a
    time (the interpreter is ~18x slower than the compiled path)."""

    def __init__(self, key: str, source_text: str, n_in: int, n_out: int, sliders=None):
        so = REF_DIR / "libshadow_eel.so"
        if not so.exists():
            raise FileNotFoundError(f"{so} (run oracle/build_ref.py where /root/reference exists)")
        # EEL2's rand() keeps ONE process-global Mersenne state (a function-static in nseel-cfunc.c), the compiled path one per
        # instance (dsp_jsfx_aot.py:5302-5305): every ShadowPlugin loads its own private copy of the library, so a script
        # that calls rand() starts from the generator's initial state like a fresh compiled instance does (render once).
        import shutil
        import tempfile
        self._tmp = tempfile.NamedTemporaryFile(prefix="libshadow_eel_", suffix=".so", delete=False)
        self._tmp.close()
        shutil.copyfile(so, self._tmp.name)
        self.lib = L = C.CDLL(self._tmp.name)
        L.zsh_new.restype = C.c_void_p
        L.zsh_new.argtypes = [C.c_char_p, C.c_void_p, C.c_void_p]
        L.zsh_error.restype = C.c_char_p
        L.zsh_error.argtypes = [C.c_void_p]
        L.zsh_free.argtypes = [C.c_void_p]
        L.zsh_prepare.argtypes = [C.c_void_p, C.c_double, C.c_void_p]
        L.zsh_process_block.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.zsh_var.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_double)]
        self.key, self.n_in, self.n_out = key, n_in, n_out
        self.sliders = sliders if sliders is not None else parse_sliders(source_text)
        self.layout = make_layout(self.sliders, lambda name: -1)
        names = (C.c_char_p * 64)(*[(s["var"].encode() if s["var"] else None) for s in self.sliders])
        self.h = L.zsh_new(source_text.encode("utf-8", "replace"), C.byref(self.layout), names)
        err = L.zsh_error(self.h).decode()
        if err:
            raise RuntimeError(f"shadow EEL2 compile of {key}: {err}")

    def render_one(self, params, x, srate=48000.0, block=512):
        """params [n_sliders] or [n_blocks][n_sliders]; x [n_in][n_frames] -> (y f32, y f64) [n_out][n_frames]."""
        x = np.ascontiguousarray(x, np.float32)
        nf = x.shape[1]
        p = np.ascontiguousarray(params, np.float32)
        p = p[None] if p.ndim == 1 else p
        assert self.lib.zsh_prepare(self.h, srate, p[0].ctypes.data) == 0, self.lib.zsh_error(self.h)
        y, yd = np.zeros((self.n_out, nf), np.float32), np.zeros((self.n_out, nf), np.float64)
        PF, PD = C.POINTER(C.c_float), C.POINTER(C.c_double)
        for b, f0 in enumerate(range(0, nf, block)):
            n = min(block, nf - f0)
            ins = (PF * max(self.n_in, 1))(*[C.cast(x[c, f0:].ctypes.data, PF) for c in range(self.n_in)])
            outs = (PF * self.n_out)(*[C.cast(y[c, f0:].ctypes.data, PF) for c in range(self.n_out)])
            outd = (PD * self.n_out)(*[C.cast(yd[c, f0:].ctypes.data, PD) for c in range(self.n_out)])
            row = p[min(b, len(p) - 1)]
            self.lib.zsh_process_block(self.h, row.ctypes.data, ins, outs, outd, self.n_in, self.n_out, n)
        return y, yd

    def var(self, name: str) -> float:
        out = C.c_double()
        return out.value if self.lib.zsh_var(self.h, name.encode(), C.byref(out)) == 0 else float("nan")

    def close(self):
        if self.h:
            self.lib.zsh_free(self.h)
            self.h = None
        try:
            os.unlink(self._tmp.name)
        except OSError:
            pass


def have_ref(key: str) -> bool:
    return (REF_DIR / f"libref_{key}.so").exists()


# ------------------------------------------------------------------------------------------------
# the plain-C restatement (oracle/za_oracle_port.c)
# ------------------------------------------------------------------------------------------------
PORT_SO = HERE / "libza_oracle_port.so"


FAUST_PORT_SO = HERE / "libza_faust_port.so"
FAUST_KEYS = ("ModTilt", "RED", "ClickBeGoneSG", "VAR", "GTS")            # Faust leaves: restatement only (parity unpinned, see za_faust_port.c)


def build_faust_port(force: bool = False) -> Path:
    import subprocess
    src = HERE / "za_faust_port.c"
    if force or not FAUST_PORT_SO.exists() or FAUST_PORT_SO.stat().st_mtime < src.stat().st_mtime:
        subprocess.run(["gcc", "-O2", "-fPIC", "-shared", "-fvisibility=hidden", "-ffp-contract=off",
                        str(src), "-o", str(FAUST_PORT_SO), "-lm", "-lpthread"], check=True)
    return FAUST_PORT_SO


def build_port(force: bool = False) -> Path:
    import subprocess
    build_faust_port(force)
    src = HERE / "za_oracle_port.c"
    if force or not PORT_SO.exists() or PORT_SO.stat().st_mtime < src.stat().st_mtime:
        subprocess.run(["gcc", "-O2", "-fPIC", "-shared", "-fvisibility=hidden", "-ffp-contract=off",
                        str(src), "-o", str(PORT_SO), "-lm", "-lpthread"], check=True)
    return PORT_SO


class PortPlugin:
    """Plain-C restatement of one plugin; same render() contract as RefPlugin."""

    CHANNELS = {"cascade8": (1, 1), "EasyExpander": (2, 2), "stft_mask2048": (2, 2), "ModTilt": (2, 2),
                "RED": (6, 6), "ClickBeGoneSG": (2, 2), "VAR": (2, 2), "GTS": (2, 2)}

    def __init__(self, key: str, sliders=None, var_index=None):
        self.key = key
        self.lib = C.CDLL(str(build_faust_port() if key in FAUST_KEYS else build_port()))
        self.fn = getattr(self.lib, f"zo_{key.lower()}_render")
        self.fn.argtypes = [C.POINTER(SliderLayout), C.c_double, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                            C.c_void_p, C.c_int, C.c_int64, C.c_int64, C.c_int, C.c_int]
        self.n_in, self.n_out = self.CHANNELS[key]
        self.sliders = sliders if sliders is not None else parse_sliders(plugin_source_text(key))
        self.layout = make_layout(self.sliders, var_index or (lambda name: -1))

    def render(self, params, x, srate=48000.0, block=512, want_f64=False, threads=0):
        x = np.ascontiguousarray(x, dtype=np.float32)
        ns, nin, nf = x.shape
        assert nin == self.n_in
        p = np.ascontiguousarray(params, dtype=np.float32)
        npb = 1 if p.ndim == 2 else p.shape[0]
        assert p.shape[-2] == ns and p.shape[-1] == len(self.sliders)
        y = np.zeros((ns, self.n_out, nf), np.float32)
        yd = np.zeros((ns, self.n_out, nf), np.float64) if want_f64 else None
        th = threads or (os.cpu_count() or 1)
        rc = self.fn(C.byref(self.layout), srate, p.ctypes.data, npb, x.ctypes.data, y.ctypes.data,
                     yd.ctypes.data if want_f64 else None, ns, nf, nf, block, th)
        assert rc == 0
        return y, yd

    def coefs(self, srate, *zone_values):
        """Slider-rate values of a Faust leaf for given zone values (tests of the product's host half)."""
        fn = getattr(self.lib, f"zo_{self.key.lower()}_coefs")
        out = np.zeros(8, np.float32)
        fn.argtypes = [C.c_double] + [C.c_float] * len(zone_values) + [C.c_void_p]
        fn(srate, *map(float, zone_values), out.ctypes.data)
        return out


def best_oracle(key: str):
    """The real reference when its build is present, else the pinned restatement."""
    return RefPlugin(key) if have_ref(key) else PortPlugin(key)
