"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol of
include/za_gpu.h, its host parameter path (slider parser + quantiser) matches the oracle's
restatement of the reference's, and compute entry points fail loudly without a GPU."""
import ctypes as C
import re
from pathlib import Path

import numpy as np
import pytest

import za_oracle as zo

REPO = Path(__file__).resolve().parent.parent


def test_exports_every_declared_symbol(built_lib):
    import za_gpu
    hdr = (REPO / "include" / "za_gpu.h").read_text()
    declared = set(re.findall(r"ZA_API\s+[\w\s\*]+?\b(za_(?:gpu|chain|batch)_\w+)\s*\(", hdr))
    assert declared == set(za_gpu.SYMBOLS), declared ^ set(za_gpu.SYMBOLS)
    L = C.CDLL(str(built_lib))
    for s in declared:
        assert hasattr(L, s), s


def test_catalogue_and_slider_layout(built_lib):
    import za_gpu
    assert "cascade8" in za_gpu.plugin_keys()
    for key in za_gpu.plugin_keys():
        info = za_gpu.plugin_info(key)
        lay = za_gpu.slider_layout(key)
        assert info.n_sliders == len(lay)
        try:
            text = zo.plugin_source_text(key)
        except KeyError:
            continue
        ours = za_gpu.parse_sliders(text)          # C++ parser on the .jsfx text
        ref = zo.parse_sliders(text)               # oracle restatement of the reference's parser
        assert len(ours) == len(ref) == len(lay)
        for a, b, c in zip(ours, ref, lay):
            assert (a.index0, bool(a.is_choice), a.alias.decode()) == (b["index0"], b["is_choice"], b["var"])
            assert (a.def_, a.min, a.max, a.step) == tuple(np.float32(b[k]) for k in ("default", "min", "max", "step"))
            assert (c.index0, c.min, c.max, c.step, c.alias) == (a.index0, a.min, a.max, a.step, a.alias)


def test_parser_grammar_cases(built_lib):
    import za_gpu
    text = "\n".join([
        "slider1:thresh_db=-40<-80,0,0.1>Threshold (dB)",
        "slider3:350<50,1200,1:sqr>Release",
        "slider8:style=1<0,4,1{Clean,Punch,Dense,Body,Wild}>--Style",
        "slider4:det_hpf_hz=0<0,20000,1:log=1000>Det HPF",
        "slider9:/some/path:none:File",
        "slider10:#name=\"x\"<string>Text",
        "slider70:1<0,1,1>ignored",
        "slider5:7<10,0>swapped range, default step",
        "slider3:1<0,1,1>duplicate index keeps the first",
    ])
    ours, ref = za_gpu.parse_sliders(text), zo.parse_sliders(text)
    assert [s.index0 for s in ours] == [r["index0"] for r in ref]
    for a, b in zip(ours, ref):
        assert (a.def_, a.min, a.max, a.step, bool(a.is_choice)) == (
            np.float32(b["default"]), np.float32(b["min"]), np.float32(b["max"]), np.float32(b["step"]), b["is_choice"])
    by = {s.index0: s for s in ours}
    assert by[7].is_choice and by[2].step == 1.0 and by[4].min == 0.0 and by[4].max == 10.0


GRAMMAR_CASES = "\n".join([
    "slider1:thresh_db=-40<-80,0,0.1>Threshold (dB)",
    "slider3:350<50,1200,1:sqr>Release",
    "slider8:style=1<0,4,1{Clean,Punch,Dense,Body,Wild}>--Style",
    "slider4:det_hpf_hz=0<0,20000,1:log=1000>Det HPF",
    "slider9:/some/path:none:File",
    "slider10:#name=\"x\"<string>Text",
    "slider11:label_only",
    "  slider 12 : a = 3 < 1 , 2 >  clamped default",
    "slider13:0.5<0,1>no step",
    "slider14:2<0,4,{A,B,C}>choices without a step",
    "slider15:1<0,3,1{ , ,}>empty choice labels",
    "slider16:5<0,10,0>zero step",
    "slider17:5<0,10,-1>negative step",
    "slider18:x=abc<0,1,0.1>unparsable default",
    "slider19:1<0,1,1:log>log without a midpoint",
    "slider70:1<0,1,1>ignored",
    "slider0:1<0,1,1>ignored too",
    "slider5:7<10,0>swapped range, default step",
    "slider3:1<0,1,1>duplicate index keeps the first",
    "// #TOOLTIP: a tooltip",
    "slider20:0<-1,1,0.01>-hidden",
    "slider21:g=-3<-80,0,0.1>G ; trailing",
])


def test_parser_equals_the_references_own_parser(built_lib):
    """za_gpu_parse_sliders against parseJsfxSliderDecls ITSELF (src/JSFXJuceProcessor.cpp:710-939, compiled from where it lies by
    oracle/build_ref.py::build_parser behind oracle/ref_parser_shim.cpp), on every .jsfx of the reference catalog (where the
    reference tree is present), on the authored workloads and on grammar corner cases.  The library lists the FLOAT
    parameters, i.e. the reference's declarations minus the string-input sliders its parameter layout skips (:2766-2772).
    The Python restatement in oracle/za_oracle.py is held to the same."""
    import glob
    import za_gpu
    if not zo.have_ref_parser():
        pytest.skip("oracle/_ref/libref_parser.so not built")
    texts = {"grammar cases": GRAMMAR_CASES}
    for f in sorted(glob.glob("/root/reference/plugins/**/*.jsfx", recursive=True) + glob.glob(str(Path(__file__).resolve().parent.parent / "zorakaudio-experimental-plugins_b200" / "plugins" / "**" / "*.jsfx"), recursive=True)):
        texts[f] = open(f, encoding="utf-8", errors="replace").read()
    n = 0
    for name, text in texts.items():
        ref = [d for d in zo.ref_parse_sliders(text) if not d["is_string"]]
        key = lambda d: (d["index0"], d["is_choice"], d["var"], np.float32(d["default"]), np.float32(d["min"]), np.float32(d["max"]), np.float32(d["step"]))
        ours = [(a.index0, bool(a.is_choice), a.alias.decode(), a.def_, a.min, a.max, a.step) for a in za_gpu.parse_sliders(text)]
        assert ours == [key(d) for d in ref], name
        assert [key(d) for d in zo.parse_sliders(text)] == [key(d) for d in ref], name
        n += len(ref)
    assert n >= 20
    by = {d["index0"]: d for d in zo.ref_parse_sliders(GRAMMAR_CASES)}
    assert by[9]["is_string"] and by[19]["hidden"] and by[7]["n_choices"] == 5 and 69 not in by       # the corner cases were exercised


def test_quantiser_matches_oracle(built_lib):
    import za_gpu
    text = zo.plugin_source_text("cascade8") + "\nslider20:mode=1<0,4,1{A,B,C,D,E}>Mode\nslider21:g=-3<-80,0,0.1>G\n"
    ours, ref = za_gpu.parse_sliders(text), zo.parse_sliders(text)
    rng = np.random.default_rng(11)
    vals = np.concatenate([rng.uniform(-100, 25000, (300, len(ref))), rng.uniform(-1, 6, (300, len(ref)))]).astype(np.float32)
    q = zo.quantise(ref, vals)
    for i in range(vals.shape[0]):
        for j in range(len(ref)):
            assert za_gpu.quantise(ours[j], float(vals[i, j])) == q[i, j]


def test_no_cpu_fallback(built_lib):
    """Without a CUDA device create() must fail loudly (ZA_ERR_CUDA), never compute on the CPU."""
    import za_gpu
    L = za_gpu.lib()
    if L.za_gpu_device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(za_gpu.ZaError, match="no CPU fallback|CUDA"):
        za_gpu.BatchProcessor("cascade8", 4)
    h = C.c_void_p()
    assert L.za_gpu_create(b"nope", 4, 48000.0, 512, 1, 0, C.byref(h)) == -2
    assert L.za_gpu_render(b"cascade8", None, None, None, 1, 1) != 0


def test_product_never_touches_oracle():
    pkg = REPO / "zorakaudio-experimental-plugins_b200"
    for p in list(pkg.rglob("*.cu")) + list(pkg.rglob("*.cpp")) + list(pkg.rglob("*.h")) + list(pkg.rglob("*.cuh")) + list(pkg.rglob("*.py")):
        t = p.read_text()
        assert "za_oracle" not in t and "oracle/" not in t and "libref_" not in t, p
