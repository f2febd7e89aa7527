"""TEST INFRASTRUCTURE ONLY -- ctypes access to the reference's own `calculate`
compiled by oracle/Makefile into oracle/_ref/ (ldscalc.h:8-65 behind
oracle/ref_wrap.cpp).  `flavour` 32 = stock float source, 64 = FP64-promoted."""
from __future__ import annotations

import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_libs: dict[int, ctypes.CDLL] = {}


def available(flavour: int = 64) -> bool:
    return os.path.exists(os.path.join(_HERE, "_ref", f"libref_ld{flavour}.so"))


def _lib(flavour: int) -> ctypes.CDLL:
    if flavour not in _libs:
        path = os.path.join(_HERE, "_ref", f"libref_ld{flavour}.so")
        lib = ctypes.CDLL(path, mode=ctypes.RTLD_LOCAL)
        lib.ref_ld_calculate.restype = ctypes.c_int
        assert lib.ref_ld_elem_bytes() == (4 if flavour == 32 else 8)
        _libs[flavour] = lib
    return _libs[flavour]


def find_openblas() -> str | None:
    """A real OpenBLAS exporting cblas_sdot, if this image has one (bundled with opencv / scipy wheels)."""
    import glob
    import site
    for sp in site.getsitepackages():
        for pat in ("opencv_python_headless.libs/libopenblas*.so*", "scipy.libs/libscipy_openblas*.so*"):
            hits = sorted(glob.glob(os.path.join(sp, pat)))
            if hits:
                return hits[0]
    import ctypes.util
    return ctypes.util.find_library("openblas")


def blas_name(flavour: int = 32) -> str:
    """Which dot the float build uses: the shim loop, or the BLAS symbol resolved through NLDSC_REF_BLAS (the variable
    is read once, at the first dot of the process)."""
    lib = _lib(flavour)
    lib.ref_shim_blas_name.restype = ctypes.c_char_p
    return lib.ref_shim_blas_name().decode()


def calculate(bedfile: str, n_snp: int, n_org: int, positions, ld_wind: float, maf: float,
              std_thr: float, rsq_thr: float, flavour: int = 64) -> dict:
    lib = _lib(flavour)
    pos = np.ascontiguousarray(positions, dtype=np.float64)
    assert pos.shape == (n_snp,)
    d = lambda: np.empty(n_snp, dtype=np.float64)
    i = lambda: np.empty(n_snp, dtype=np.int32)
    out = {"l2": d(), "l2d": d(), "maf": d(), "residuals_std": d(),
           "l2_ws": i(), "l2d_ws": i(), "l2d_wse": i()}
    err = ctypes.create_string_buffer(512)
    P = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    rc = lib.ref_ld_calculate(os.fsencode(bedfile), ctypes.c_int(n_snp), ctypes.c_int(n_org),
                              ctypes.c_double(ld_wind), ctypes.c_double(maf),
                              ctypes.c_double(std_thr), ctypes.c_double(rsq_thr), P(pos),
                              P(out["l2"]), P(out["l2d"]), P(out["maf"]), P(out["residuals_std"]),
                              P(out["l2_ws"]), P(out["l2d_ws"]), P(out["l2d_wse"]),
                              err, ctypes.c_size_t(512))
    if rc == 1:
        raise ValueError(err.value.decode())
    if rc != 0:
        raise RuntimeError(err.value.decode())
    return out
