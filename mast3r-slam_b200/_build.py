"""In-tree build of the CUDA extension: nvcc -> mast3r-slam_b200/libm3s_b200.so (sm_100a only).

The shared library has a pure C ABI (include/m3s.h) and no torch dependency; it travels to the GPU
box with the repo snapshot.  match_exact.cu is compiled with -fmad=false on top of its explicit
round-to-nearest intrinsics (bit-exact iter_proj)."""
from __future__ import annotations

import os
import subprocess

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
SO = os.path.join(PKG, "libm3s_b200.so")
OBJ_DIR = os.path.join(PKG, "build")

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden",
          "-I", os.path.join(ROOT, "include"), "-I", CSRC, "-ccbin", "g++"]
SOURCES = {
    "api.cu": [],
    "match_exact.cu": ["-fmad=false"],
    "refine.cu": [],
    "gn.cu": [],
    "graph_host.cu": ["-fmad=false"],
    "host_pipeline.cu": [],
    "track.cu": ["-fmad=false"],
    "metal_mode.cu": ["-fmad=false"],
}


def _newer(target, deps):
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(d) <= t for d in deps)


def build_variant(name: str, defines: list[str]) -> str:
    """Development helper: build libm3s_b200_<name>.so with extra -D flags (A/B experiments;
    select it at run time with M3S_LIB=<path>)."""
    odir = os.path.join(OBJ_DIR, name)
    os.makedirs(odir, exist_ok=True)
    objs = []
    for src, extra in SOURCES.items():
        o = os.path.join(odir, src.replace(".cu", ".o"))
        subprocess.check_call(["nvcc", *ARCH, *COMMON, *extra, *[f"-D{d}" for d in defines], "-c",
                               os.path.join(CSRC, src), "-o", o])
        objs.append(o)
    so = os.path.join(PKG, f"libm3s_b200_{name}.so")
    subprocess.check_call(["nvcc", *ARCH, "-shared", "-o", so, *objs, "-ccbin", "g++"])
    return so


def build_native(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJ_DIR, exist_ok=True)
    headers = [os.path.join(CSRC, "m3s_common.cuh"), os.path.join(ROOT, "include", "m3s.h")]
    objs = []
    rebuilt = False
    for src, extra in SOURCES.items():
        s = os.path.join(CSRC, src)
        o = os.path.join(OBJ_DIR, src.replace(".cu", ".o"))
        objs.append(o)
        if force or not _newer(o, [s] + headers):
            cmd = ["nvcc", *ARCH, *COMMON, *extra, "-c", s, "-o", o]
            if verbose:
                cmd.insert(1, "-Xptxas=-v")
            subprocess.check_call(cmd)
            rebuilt = True
    if rebuilt or not os.path.exists(SO):
        subprocess.check_call(["nvcc", *ARCH, "-shared", "-o", SO, *objs, "-ccbin", "g++"])
    return SO


if __name__ == "__main__":
    import sys
    print(build_native(force="--force" in sys.argv, verbose="-v" in sys.argv))
