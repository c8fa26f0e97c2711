"""Build the in-tree CUDA library (sm_100a only) and the LINKS_CPP host binary.

    python -m links_b200.build          # liblinks_gpu.so + bin/LINKS_CPP
    python -m links_b200.build --force

Artifacts are git-ignored but travel to the GPU box with the repo snapshot.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "liblinks_gpu.so")
HOST_BIN = os.path.join(HERE, "bin", "LINKS_CPP")
LAYOUT_BIN = os.path.join(HERE, "bin", "LINKS_layout")  # the Perl stage (bin/LINKS.pl) in C++; also installed as bin/LINKS.pl

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC,-Wall",
]


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: cannot build liblinks_gpu.so (there is no CPU fallback)")


def _newer(target: str, sources: list[str]) -> bool:
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(s) <= t for s in sources)


def _sources(exts: tuple[str, ...]) -> list[str]:
    out = []
    for d, _, files in os.walk(CSRC):
        out += [os.path.join(d, f) for f in files if f.endswith(exts)]
    out.append(os.path.join(ROOT, "include", "links_gpu.h"))
    return out


def build_lib(force: bool = False, verbose: bool = False) -> str:
    srcs = _sources((".cu", ".cuh", ".h"))
    if not force and _newer(LIB, srcs):
        return LIB
    cmd = [_nvcc(), *NVCC_FLAGS, "-shared", "-o", LIB, os.path.join(CSRC, "links_gpu.cu")]
    cmd[1:1] = os.environ.get("LINKS_NVCC_DEFS", "").split()  # kernel-variant experiments: -DLINKS_EXTRACT_TILE=256 ...
    if verbose:
        cmd[1:1] = ["-Xptxas", "-v"]
        print(" ".join(cmd))
    subprocess.run(cmd, check=True)
    return LIB


def build_host(force: bool = False) -> str | None:
    src = os.path.join(CSRC, "host", "links_cpp_main.cpp")
    if not os.path.exists(src):
        return None
    srcs = _sources((".cpp", ".hpp", ".h"))
    if not force and _newer(HOST_BIN, srcs + [LIB]):
        return HOST_BIN
    os.makedirs(os.path.dirname(HOST_BIN), exist_ok=True)
    cmd = ["g++", "-std=c++17", "-O2", "-Wall", "-pthread", "-I", os.path.join(ROOT, "include"), src,
           "-o", HOST_BIN, "-L", HERE, "-llinks_gpu", "-Wl,-rpath,$ORIGIN/.."]
    subprocess.run(cmd, check=True)
    return HOST_BIN


def build_layout(force: bool = False) -> str:
    """CPU only (no CUDA): usable on any machine, also by the `-m "not gpu"` tests."""
    src = os.path.join(CSRC, "host", "links_layout_main.cpp")
    srcs = [src, os.path.join(CSRC, "host", "links_layout.hpp")]
    if force or not _newer(LAYOUT_BIN, srcs):
        os.makedirs(os.path.dirname(LAYOUT_BIN), exist_ok=True)
        subprocess.run(["g++", "-std=c++17", "-O2", "-Wall", src, "-o", LAYOUT_BIN], check=True)
    alias = os.path.join(os.path.dirname(LAYOUT_BIN), "LINKS.pl")  # what the unmodified LINKS-make calls (bin/LINKS-make:124)
    if force or not _newer(alias, [LAYOUT_BIN]):
        shutil.copy2(LAYOUT_BIN, alias)
    return LAYOUT_BIN


SYNTHGEN_DIR = os.path.join(ROOT, "tools", "synthgen")
SYNTHGEN_LIB = os.path.join(SYNTHGEN_DIR, "libsynth_gpu.so")


def build_synthgen(force: bool = False) -> str:
    """Device-side generator of the human-scale synthetic workload (bench / test infrastructure, its own library)."""
    src = os.path.join(SYNTHGEN_DIR, "synth_gen.cu")
    if force or not _newer(SYNTHGEN_LIB, [src]):
        subprocess.run([_nvcc(), *NVCC_FLAGS, "-shared", "-o", SYNTHGEN_LIB, src], check=True)
    return SYNTHGEN_LIB


GATHER_PEAK_BIN = os.path.join(ROOT, "tools", "gather_peak")


def build_tools(force: bool = False) -> None:
    """Measurement probes under tools/ (random-sector gather ceiling, random-word atomic ceiling, DRAM bytes per load flavour)."""
    for name in ("gather_peak", "atomic_peak", "load_variants"):
        src, dst = os.path.join(ROOT, "tools", name + ".cu"), os.path.join(ROOT, "tools", name)
        if os.path.exists(src) and (force or not _newer(dst, [src])):
            subprocess.run([_nvcc(), *NVCC_FLAGS, "-o", dst, src], check=True)


def build_all(force: bool = False, verbose: bool = False) -> None:
    build_lib(force, verbose)
    build_synthgen(force)
    build_tools(force)
    build_host(force)
    build_layout(force)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(LIB)
