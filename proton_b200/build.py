"""Builds libproton_em.so (and the EMtool_analytical-compatible launcher) in-tree with nvcc for sm_100a."""
from __future__ import annotations

import os
import shutil
import subprocess
from pathlib import Path

PKG = Path(__file__).resolve().parent
CSRC = PKG / "csrc"
LIB = PKG / "libproton_em.so"
LAUNCHER = PKG.parent / "bin" / "EMtool_analytical"

# -fmad=false: no implicit contraction in device code — every fused multiply-add of the kernels is written as fma() in
# the source, which is what lets the CPU twin of the tests repeat the arithmetic bit for bit (82 of 38 700 DFMAs were implicit).
NVCC_FLAGS = ["-std=c++17", "-O3", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-fmad=false",
              "-Xcompiler", "-fPIC,-ffp-contract=off"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found")


def _newer(target: Path, sources) -> bool:
    if not target.exists():
        return False
    t = target.stat().st_mtime
    return all(Path(s).stat().st_mtime <= t for s in sources)


def build_library(force: bool = False, verbose: bool = False) -> Path:
    sources = sorted(list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cpp")))
    sources = [s for s in sources if s.name != "emtool_main.cpp"]
    deps = sources + list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.inc")) + [PKG.parent / "include" / "proton_em.h"]
    if not force and _newer(LIB, deps):
        return LIB
    extra = os.environ.get("PROTON_EM_NVCC_EXTRA", "").split()
    cmd = [_nvcc(), *NVCC_FLAGS, *extra, "-shared", "-o", str(LIB), *[str(s) for s in sources]]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stderr)
    return LIB


def build_launcher(force: bool = False) -> Path:
    main = CSRC / "emtool_main.cpp"
    if not main.exists():
        return LAUNCHER
    build_library(force=force)
    if not force and _newer(LAUNCHER, [main, LIB]):
        return LAUNCHER
    LAUNCHER.parent.mkdir(parents=True, exist_ok=True)
    cmd = [_nvcc(), "-std=c++17", "-O2", "-o", str(LAUNCHER), str(main), f"-L{PKG}", "-lproton_em",
           "-Xlinker", "-rpath", "-Xlinker", "$ORIGIN/../proton_b200"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("launcher build failed:\n" + res.stdout + res.stderr)
    return LAUNCHER


if __name__ == "__main__":
    print(build_library(force=True, verbose=True))
    print(build_launcher(force=True))
