"""Builds tinypathtracer_b200/lib/libtpt_b200.so with nvcc for sm_100a (cross-compiles without a GPU)."""
from __future__ import annotations

import os
import subprocess
import sys
from pathlib import Path

PKG = Path(__file__).resolve().parent
ROOT = PKG.parent
CSRC = PKG / "csrc"
LIB = PKG / "lib" / "libtpt_b200.so"
SOURCES = ["capi.cu", "bvh.cu", "render.cu", "env.cu", "comm.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC",
              "-ccbin", "/usr/bin/g++", f"-I{ROOT / 'include'}", "--expt-relaxed-constexpr"]


def _stale() -> bool:
    if not LIB.exists():
        return True
    t = LIB.stat().st_mtime
    deps = list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cuh")) + list((ROOT / "include").glob("*.h"))
    return any(d.stat().st_mtime > t for d in deps)


def build_library(force: bool = False, verbose: bool = False, defines=(), out: Path | None = None) -> Path:
    """`defines`/`out` build a tuning variant beside the product library (loaded with TPT_B200_LIB=<path>)."""
    if out is None and not force and not _stale():
        return LIB
    out = out or LIB
    out.parent.mkdir(parents=True, exist_ok=True)
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc, *NVCC_FLAGS, *[f"-D{d}" for d in defines], "-shared", "-o", str(out), *[str(CSRC / s) for s in SOURCES]]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("nvcc failed building libtpt_b200.so")
    if verbose:
        print(r.stderr)
    return out


HOST_LIB = PKG / "lib" / "libtpt_host.so"
HOST_EXE = PKG / "lib" / "tpt_render"
HOST_SOURCES = ["host.cpp", "image_io.cpp", "path_tracer.cpp"]


def build_host(force: bool = False) -> Path:
    """Host C++ layer (include/tpt/host.hpp): libtpt_host.so + the tpt_render CLI, linked against libtpt_b200.so.
    -ffp-contract=off: the TRS matrices must equal the reference's bit for bit."""
    srcs = [CSRC / "host" / s for s in HOST_SOURCES]
    deps = srcs + [CSRC / "host" / "main.cpp"] + list((ROOT / "include" / "tpt").glob("*.hpp")) + [ROOT / "include" / "tpt.h", LIB]
    if not force and HOST_LIB.exists() and HOST_EXE.exists() and all(d.stat().st_mtime <= min(HOST_LIB.stat().st_mtime, HOST_EXE.stat().st_mtime) for d in deps):
        return HOST_LIB
    gxx = os.environ.get("CXX", "/usr/bin/g++")
    common = [gxx, "-O2", "-std=c++17", "-fPIC", "-ffp-contract=off", "-Wall", f"-I{ROOT / 'include'}"]
    link = [f"-L{LIB.parent}", "-ltpt_b200", "-lz", "-Wl,-rpath,$ORIGIN"]
    for cmd in ([*common, "-shared", "-o", str(HOST_LIB), *map(str, srcs), *link],
                [*common, "-o", str(HOST_EXE), str(CSRC / "host" / "main.cpp"), f"-L{LIB.parent}", "-ltpt_host", *link]):
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            sys.stderr.write(r.stdout + r.stderr)
            raise RuntimeError("g++ failed building the host layer")
    return HOST_LIB


if __name__ == "__main__":
    defs = [a[2:] for a in sys.argv[1:] if a.startswith("-D")]
    outs = [a[2:] for a in sys.argv[1:] if a.startswith("-o")]
    print(build_library(force="--force" in sys.argv, verbose="-v" in sys.argv, defines=defs, out=Path(outs[0]).resolve() if outs else None))
    if not outs:
        print(build_host(force="--force" in sys.argv))
