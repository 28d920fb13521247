"""Build libza_gpu.so (C-ABI + sm_100a kernels) in-tree with nvcc.  No JIT cache, no torch."""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG = Path(__file__).resolve().parent.parent
CSRC = PKG / "csrc"
BUILD = PKG / "build"
LIB = PKG / "libza_gpu.so"

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC,-fvisibility=hidden,-ffp-contract=off",
          "-ftz=false", "-prec-div=true", "-prec-sqrt=true"]


# Faust leaves restate the .dsp's separate float multiplies and adds: no FMA contraction in those translation units
PER_FILE = {"p_faust.cu": ["-fmad=false"], "k_gts.cu": ["-fmad=false"]}
if os.environ.get("ZA_EE_KQ"):          # measurement builds of the EasyExpander v3 kernel: streams per helper warp (2 or 4)
    PER_FILE["k_easyexpander_v3.cu"] = ["-DZA_EE_KQ=" + os.environ["ZA_EE_KQ"]]
if os.environ.get("ZA_EE_PPL"):         # ... and band pairs per band-warp lane (1 or 2)
    PER_FILE.setdefault("k_easyexpander_v3.cu", []).append("-DZA_EE_PPL=" + os.environ["ZA_EE_PPL"])
if os.environ.get("ZA_STAGED_CTAS"):    # measurement builds of the stage pipelines: CTAs per SM they are compiled for, shared-memory budget in KB
    for _f in ("g_RTT.cu", "g_SOMA.cu", "g_ATTACK.cu"):
        PER_FILE[_f] = ["-DZA_STAGED_MIN_CTAS(K)=" + os.environ["ZA_STAGED_CTAS"], "-DZA_STAGED_SMEM_KB=" + os.environ.get("ZA_STAGED_SMEM_KB", "110")]


def sources():
    return sorted(list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cpp")) + list((CSRC / "gen").glob("*.cu")))


def _stale(obj: Path, src: Path) -> bool:
    if not obj.exists():
        return True
    t = obj.stat().st_mtime
    deps = [src] + list(CSRC.glob("*.h")) + list(CSRC.glob("*.cuh")) + list((CSRC / "gen").glob("*.h")) + [PKG.parent / "include" / "za_gpu.h"]
    return any(d.stat().st_mtime > t for d in deps)


def build_lib(force: bool = False, verbose: bool = False) -> Path:
    BUILD.mkdir(exist_ok=True)
    objs, jobs = [], []
    for src in sources():
        obj = BUILD / (src.stem + ".o")
        objs.append(obj)
        if force or _stale(obj, src):
            cmd = [NVCC, *ARCH, *COMMON, *PER_FILE.get(src.name, []), "-x", "cu", "-c", str(src), "-o", str(obj)]
            if verbose:
                cmd.insert(1, "-Xptxas=-v")
            jobs.append(cmd)

    def run(cmd):
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + r.stdout + r.stderr)
        return r.stdout + r.stderr

    with ThreadPoolExecutor(max_workers=4) as ex:
        logs = list(ex.map(run, jobs))
    if verbose:
        for l in logs:
            print(l)
    if jobs or not LIB.exists():
        run([NVCC, *ARCH, "-shared", "-o", str(LIB), *map(str, objs), "-Xcompiler", "-fPIC", "-cudart", "static"])
    build_bounce()
    return LIB


BOUNCE = PKG / "za_bounce"


def build_bounce(force: bool = False) -> Path:
    """The offline bounce driver (host/za_bounce.cpp): plain C++ over the C ABI, linked against the in-tree library."""
    srcs = [PKG / "host" / "za_bounce.cpp", PKG / "host" / "za_wav.h", PKG.parent / "include" / "za_gpu.h"]
    if force or not BOUNCE.exists() or any(x.stat().st_mtime > BOUNCE.stat().st_mtime for x in srcs + [LIB]):
        r = subprocess.run([os.environ.get("CXX", "g++"), "-O2", "-std=c++17", "-Wall", str(srcs[0]), "-o", str(BOUNCE),
                            "-L" + str(PKG), "-lza_gpu", "-Wl,-rpath,$ORIGIN"], capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("g++ failed:\n" + r.stdout + r.stderr)
    return BOUNCE


if __name__ == "__main__":
    print(build_lib(force="--force" in sys.argv, verbose="-v" in sys.argv))
