#!/usr/bin/env python3
"""oracle/build_ref.py -- TEST INFRASTRUCTURE: build the real reference into oracle/_ref/.

For every plugin in PLUGINS this runs the reference's OWN compiler, read-only, from where it lies
(`/root/reference/dsp_jsfx_aot.py --opt 2`, the same invocation as scripts/build.py:385-393) on the
plugin's .jsfx source, and links the emitted object with oracle/ref_shim.c (and, for FFT users,
oracle/ref_fft_hooks.c over /root/reference/src/WDL/fft.c compiled in place) into

    oracle/_ref/libref_<key>.so   + <key>.json (the compiler's own --meta: var table, channel counts)

Nothing from /root/reference is copied into the repo: sources are compiled where they lie and only
build outputs land in oracle/_ref/ (git-ignored, but shipped to the GPU box by gpurun).
On a machine without /root/reference (the GPU box) this script is a no-op and the prebuilt files
are used.
"""
from __future__ import annotations

import os
import subprocess
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
REPO = HERE.parent
REF = Path(os.environ.get("ZA_REFERENCE_ROOT", "/root/reference"))
OUT = HERE / "_ref"
PKG = REPO / "zorakaudio-experimental-plugins_b200"

# key -> (jsfx source, needs FFT hooks)
PLUGINS = {
    "cascade8": (PKG / "plugins/Workloads/cascade8/src/cascade8.jsfx", False),
    "stft_mask2048": (PKG / "plugins/Workloads/stft_mask2048/src/stft_mask2048.jsfx", True),
    "decrust512": (PKG / "plugins/Workloads/decrust512/src/decrust512.jsfx", True),
    "fftprobe": (PKG / "plugins/Workloads/fftprobe/src/fftprobe.jsfx", True),
    "semprobe": (PKG / "plugins/Workloads/semprobe/src/semprobe.jsfx", False),
    "Alias": (REF / "plugins/Spectral/Alias/src/Alias.jsfx", False),
    "EasyExpander": (REF / "plugins/Dynamics/EasyExpander/src/EasyExpander.jsfx", False),
    "RTT": (REF / "plugins/Dynamics/RTT/src/Reverb Tail Tamer.jsfx", False),
    "SOMA": (REF / "plugins/Dynamics/SOMA/src/SOMA.jsfx", False),
    "ATTACK": (REF / "plugins/Dynamics/ATTACK/src/ATTACK.jsfx", False),
    "ERBTilt": (REF / "plugins/Spectral/ERBTilt/src/ERB Tilt.jsfx", False),
    "SpectralStabilizer": (REF / "plugins/Spectral/SpectralStabilizer/src/Spectral Stabilizer.jsfx", False),
    "DDT": (REF / "plugins/Spatialization/DDT/src/DDT.jsfx", False),
    "DPT": (REF / "plugins/Spatialization/DPT/src/DPT.jsfx", False),
    "SaliencePush": (REF / "plugins/Spatialization/SaliencePush/src/SP.jsfx", False),
    "ADS": (REF / "plugins/Ambience/ADS/src/ADS.jsfx", False),
    "BedRock": (REF / "plugins/Spectral/BedRock/src/BedRock.jsfx", False),
    "Roomalizer": (REF / "plugins/Spatialization/Roomalizer/src/Roomalizer.jsfx", False),
    "TSEQ": (REF / "plugins/Spectral/TSEQ/src/TSEQ.jsfx", False),
}


def _newer(dst: Path, *srcs: Path) -> bool:
    if not dst.exists():
        return False
    t = dst.stat().st_mtime
    return all(s.exists() and s.stat().st_mtime <= t for s in srcs)


def build_one(key: str, force: bool = False) -> Path | None:
    src, needs_fft = PLUGINS[key]
    so = OUT / f"libref_{key}.so"
    if not REF.exists() or not src.exists():
        return so if so.exists() else None
    shim = HERE / "ref_shim.c"
    hooks = HERE / "ref_fft_hooks.c"
    deps = [src, shim] + ([hooks] if needs_fft else [])
    if not force and _newer(so, *deps):
        return so
    OUT.mkdir(exist_ok=True)
    obj, hdr, meta = OUT / f"{key}.o", OUT / f"{key}.h", OUT / f"{key}.json"
    subprocess.run(
        [sys.executable, str(REF / "dsp_jsfx_aot.py"), str(src), "--out-obj", str(obj), "--out-h", str(hdr),
         "--meta", str(meta), "--out-ll", os.devnull, "--opt", "2"],
        check=True, stdout=subprocess.DEVNULL)
    cmd = ["gcc", "-O2", "-fPIC", "-shared", "-fvisibility=hidden", "-ffp-contract=off",
           f'-DZREF_HEADER="{hdr}"', str(shim), str(obj)]
    if needs_fft:
        # WDL_FFT_REALSIZE=8: cmake/plugin/CMakeLists.txt:224
        cmd += ["-DWDL_FFT_REALSIZE=8", f"-I{REF / 'src'}", f'-DZREF_WDL_FFT_C="{REF / "src/WDL/fft.c"}"', str(hooks)]
    cmd += ["-o", str(so), "-lm", "-lpthread"]
    subprocess.run(cmd, check=True)
    return so


def build_shadow(force: bool = False) -> Path | None:
    """The secondary oracle: oracle/shadow_eel.cpp over the reference's own WDL EEL2 interpreter (portable target), compiled
    from /root/reference/src/WDL where it lies into oracle/_ref/libshadow_eel.so."""
    so = OUT / "libshadow_eel.so"
    src = HERE / "shadow_eel.cpp"
    wdl = REF / "src" / "WDL"
    if not REF.exists() or not (wdl / "eel2" / "nseel-compiler.c").exists():
        return so if so.exists() else None
    if not force and _newer(so, src):
        return so
    OUT.mkdir(exist_ok=True)
    objs = []
    flags = ["-O2", "-fPIC", "-DEEL_TARGET_PORTABLE=1", "-DWDL_FFT_REALSIZE=8", f"-I{REF / 'src'}", "-w"]
    for name in ("nseel-compiler", "nseel-eval", "nseel-lextab", "nseel-ram", "nseel-yylex", "nseel-cfunc"):
        o = OUT / f"eel2_{name}.o"
        subprocess.run(["gcc", *flags, "-c", str(wdl / "eel2" / f"{name}.c"), "-o", str(o)], check=True)
        objs.append(str(o))
    o = OUT / "eel2_fft.o"
    subprocess.run(["gcc", *flags, "-c", str(wdl / "fft.c"), "-o", str(o)], check=True)
    objs.append(str(o))
    subprocess.run(["g++", *flags, "-std=c++17", "-fvisibility=hidden", "-shared", str(src), *objs, "-o", str(so), "-lm"], check=True)
    return so


def _cxx_block(text: str, start: int) -> str:
    """The declaration that begins at `start`, up to and including its closing brace (plus a `;` after a struct).  A small
    C++ scanner: braces inside string / raw-string / char literals and comments do not count."""
    i, depth, n = start, 0, len(text)
    while i < n:
        c = text[i]
        if text.startswith("//", i):
            i = text.index("\n", i)
        elif text.startswith("/*", i):
            i = text.index("*/", i) + 2
        elif text.startswith('R"(', i):
            i = text.index(')"', i) + 2
        elif c == '"':
            i += 1
            while text[i] != '"':
                i += 2 if text[i] == "\\" else 1
            i += 1
        elif c == "'":
            i += 1
            while text[i] != "'":
                i += 2 if text[i] == "\\" else 1
            i += 1
        else:
            if c == "{":
                depth += 1
            elif c == "}":
                depth -= 1
                if depth == 0:
                    i += 1
                    while i < n and text[i] in " \t":
                        i += 1
                    return text[start:i + 1] if i < n and text[i] == ";" else text[start:i]
            i += 1
    raise RuntimeError("unbalanced braces in the reference source")


PARSER_DECLS = ["struct JsfxSliderDecl", "static inline std::string trimAscii (std::string s)",
                "static std::vector<std::string> splitTopLevelCommas (const std::string& s)",
                "static bool parseFloat (const std::string& s, float& out)",
                "static juce::String parseJsfxStringDefaultToken (std::string tok)",
                "static std::vector<JsfxSliderDecl> parseJsfxSliderDecls (const char* jsfxText, juce::String* outHelpText = nullptr)"]


def build_parser(force: bool = False) -> Path | None:
    """The reference's own slider-declaration parser (src/JSFXJuceProcessor.cpp: JsfxSliderDecl :526, trimAscii :567,
    splitTopLevelCommas / parseFloat / parseJsfxStringDefaultToken :633-704, parseJsfxSliderDecls :710-939) behind a C entry
    point: the declarations are cut out of the source where it lies into a temporary oracle/_ref/ref_parser_extract.inc, compiled
    inside oracle/ref_parser_shim.cpp (which supplies the three JUCE names they use), and the extract is deleted again."""
    src = REF / "src" / "JSFXJuceProcessor.cpp"
    lib = OUT / "libref_parser.so"
    if not src.exists():
        return lib if lib.exists() else None
    shim = HERE / "ref_parser_shim.cpp"
    if lib.exists() and not force and not _newer(lib, src, shim, Path(__file__)):
        return lib
    OUT.mkdir(exist_ok=True)
    text = src.read_text(errors="replace")
    parts = []
    for d in PARSER_DECLS:
        at = text.find(d)
        if at < 0:
            raise RuntimeError(f"reference parser: `{d}` not found in {src}")
        parts.append(f"// ---- {src.name}, line {text.count(chr(10), 0, at) + 1} ----\n" + _cxx_block(text, at) + "\n")
    (OUT / "ref_parser_extract.inc").write_text("// generated by oracle/build_ref.py::build_parser from the reference source; not tracked\n" + "\n".join(parts))
    r = subprocess.run([os.environ.get("CXX", "g++"), "-O1", "-std=c++17", "-shared", "-fPIC", "-fvisibility=hidden", "-static-libstdc++", "-static-libgcc", "-Wl,-Bsymbolic", "-Wl,--exclude-libs,ALL", "-I", str(HERE),
                        str(shim), "-o", str(lib)], capture_output=True, text=True)
    (OUT / "ref_parser_extract.inc").unlink()           # only the compiled library stays (and travels to the GPU box)
    if r.returncode != 0:
        raise RuntimeError("reference parser build failed:\n" + r.stdout + r.stderr)
    return lib


def build_all(keys=None, force: bool = False) -> dict:
    res = {}
    if keys is None:
        try:
            res["shadow_eel"] = build_shadow(force)
        except subprocess.CalledProcessError as e:
            print(f"[build_ref] shadow_eel: FAILED ({e})", file=sys.stderr)
            res["shadow_eel"] = None
        try:
            res["ref_parser"] = build_parser(force)
        except (RuntimeError, ValueError) as e:
            print(f"[build_ref] ref_parser: FAILED ({e})", file=sys.stderr)
            res["ref_parser"] = None
    for k in (keys or PLUGINS):
        try:
            res[k] = build_one(k, force)
        except subprocess.CalledProcessError as e:  # keep going: one broken leaf must not hide the others
            print(f"[build_ref] {k}: FAILED ({e})", file=sys.stderr)
            res[k] = None
    return res


if __name__ == "__main__":
    keys = [a for a in sys.argv[1:] if not a.startswith("-")] or None
    for k, v in build_all(keys, force="--force" in sys.argv).items():
        print(f"{k:22s} {v}")
