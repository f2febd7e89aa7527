"""In-tree build of the native code (sm_100a only).

  nldsc_b200/libnldsc_ld.so                 CUDA kernels + C ABI (include/nldsc_ld.h)
  nldsc_b200/_ldscore.<abi>.so              pybind11 shim with the reference's `_ldscore` surface

Plays the role of the reference's nldsc/ldscore/_ldscore/CMakeLists.txt + Makefile:5-8
(build the extension, drop the .so next to the Python package).  nvcc cross-compiles without
a GPU; the .so files are git-ignored but travel to the GPU box with the tree.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libnldsc_ld.so")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC"]
# unpack.cu holds the float MAF path and the closed-form per-SNP statistics: no FMA contraction there
SOURCES = {"unpack.cu": ["-fmad=false"], "plan.cu": ["-fmad=false"], "band_tc.cu": [], "band_simt.cu": [], "session.cu": [],
           "static_stats.cu": []}
HEADERS = ["common.cuh", "epilogue.cuh", "session.cuh", os.path.join("..", "..", "include", "nldsc_ld.h")]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _stale(target: str, deps: list[str]) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _run(cmd: list[str]) -> None:
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("build failed: " + " ".join(cmd) + "\n" + r.stdout + r.stderr)


def ext_path() -> str:
    return os.path.join(HERE, "_ldscore" + sysconfig.get_config_var("EXT_SUFFIX"))


def build_variant(tag: str, defines: list[str]) -> str:
    """An A/B build of the same sources with extra -D flags -> nldsc_b200/libnldsc_ld_<tag>.so (select it with NLDSC_LIB).
    Measurement only: never loaded unless asked for."""
    obj = os.path.join(OBJ, tag)
    os.makedirs(obj, exist_ok=True)
    nvcc = _nvcc()
    objs = []
    for src, extra in SOURCES.items():
        o = os.path.join(obj, src.replace(".cu", ".o"))
        objs.append(o)
        _run([nvcc] + ARCH + COMMON + extra + ["-D" + d for d in defines] + ["-c", os.path.join(CSRC, src), "-o", o])
    out = os.path.join(HERE, f"libnldsc_ld_{tag}.so")
    _run([nvcc] + ARCH + ["-shared", "-o", out] + objs)
    return out


def build_native(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJ, exist_ok=True)
    nvcc = _nvcc()
    hdrs = [os.path.join(CSRC, h) for h in HEADERS] + [os.path.abspath(__file__)]
    objs = []
    for src, extra in SOURCES.items():
        s = os.path.join(CSRC, src)
        o = os.path.join(OBJ, src.replace(".cu", ".o"))
        objs.append(o)
        if force or _stale(o, [s] + hdrs):
            if verbose:
                print("nvcc", src, file=sys.stderr)
            _run([nvcc] + ARCH + COMMON + extra + ["-c", s, "-o", o])
    if force or _stale(LIB, objs):
        _run([nvcc] + ARCH + ["-shared", "-o", LIB] + objs)
    return LIB


def build_module(force: bool = False) -> str:
    """The pybind11 `_ldscore` extension, linked against libnldsc_ld.so (rpath $ORIGIN)."""
    import pybind11
    build_native(force)
    out = ext_path()
    src = os.path.join(CSRC, "ldscore_module.cpp")
    if force or _stale(out, [src, LIB, os.path.join(HERE, "..", "include", "nldsc_ld.h")]):
        cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
        _run([cxx, "-O2", "-std=c++17", "-shared", "-fPIC", "-fvisibility=hidden",
              "-I" + pybind11.get_include(), "-I" + sysconfig.get_paths()["include"], src, "-o", out,
              "-L" + HERE, "-lnldsc_ld", "-Wl,-rpath,$ORIGIN"])
    return out


if __name__ == "__main__":
    build_native(force="--force" in sys.argv, verbose=True)
    print(build_module(force="--force" in sys.argv))
