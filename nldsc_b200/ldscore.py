"""Host-side mirror of the reference's `ld` Python driver, on top of the B200 `_ldscore` module.

The reference's own Python layer (nldsc/ldscore/routine.py, common.py, nldsc/__main__.py) is the
caller of the hot path and stays unchanged when the `_ldscore` extension is swapped (see
INTEGRATION.md).  This module restates just enough of it, under the same names and with the same
argument meaning and error behaviour, that the path can be driven and tested end to end without
the reference tree: `python -m nldsc_b200 ld --bfile ... --ld-wind-cm ... -maf ... --std-thr ...`.

  estimate_lds                   ldscore/routine.py:51-102
  make_output (column layout)    ldscore/routine.py:32-48
  window / threshold validation  ldscore/common.py:10-36,146-182
  .bim / .fam parsing            ldscore/common.py:86-89,114-117,130-133
"""
from __future__ import annotations

import os

import numpy as np
import pandas as pd

BIM_COLUMNS = ("CHR", "SNP", "CM", "BP", "A1", "A2")


class NLDSCParameterError(ValueError):
    """Same role as the reference's core.common.NLDSCParameterError."""


def _window(ld_wind, metric: str) -> tuple[float, str]:
    w = float(ld_wind)
    if metric == "kbp":
        w, metric = w * 1000.0, "bp"
    if metric not in ("bp", "cm"):
        raise NLDSCParameterError("Invalid metric")
    if w <= 0:
        raise NLDSCParameterError("The ld-window must be greater than 0")
    if metric == "bp" and w > 5e6:
        raise NLDSCParameterError("The ld-window cannot be larger than 5 Mbp")
    if metric == "cm" and w > 100:
        raise NLDSCParameterError("The ld-window cannot be larger than 100 cm")
    return w, metric


def _unit_interval(value, hi: float, what: str) -> float:
    v = float(value)
    if not (0 <= v < hi):
        raise NLDSCParameterError(f"{what} must be between 0 and {hi:g}!")
    return v


def _bfile_paths(bfile: str) -> tuple[str, str, str]:
    path = os.path.abspath(bfile)
    root, ext = os.path.splitext(path)
    if ext in (".bed", ".bim", ".fam"):
        path = root
    paths = tuple(path + e for e in (".bed", ".bim", ".fam"))
    for p in paths:
        if not os.path.exists(p):
            raise FileNotFoundError(f'No such file: "{p}"')
    return paths


def _module():
    try:
        from . import _ldscore
    except ImportError as e:          # same contract as the reference's routines.py:5-10 stub
        raise RuntimeError("nldsc_b200._ldscore is not built: run `python -m nldsc_b200.build` "
                           "(there is no CPU fallback)") from e
    return _ldscore


def make_output(bim: pd.DataFrame, ld, *, extra: bool = False) -> pd.DataFrame:
    """CHR SNP BP L2 L2D [MAF WSA WSD WSDE RSTD] -- the layout h2's LDScoreReader reads by name."""
    out = pd.DataFrame({"CHR": bim["CHR"], "SNP": bim["SNP"], "BP": bim["BP"],
                        "L2": np.asarray(ld.l2, dtype=np.float64), "L2D": np.asarray(ld.l2d, dtype=np.float64)})
    if extra:
        out["MAF"] = np.asarray(ld.maf, dtype=np.float64)
        out["WSA"] = np.asarray(ld.l2_ws, dtype=np.int64)
        out["WSD"] = np.asarray(ld.l2d_ws, dtype=np.int64)
        out["WSDE"] = np.asarray(ld.l2d_wse, dtype=np.int64)
        out["RSTD"] = np.asarray(ld.residuals_std, dtype=np.float64)
    return out


def estimate_lds(bfile: str, ld_wind: float, wind_metric: str, maf_thr: float = 1e-5, std_thr: float = 1e-5,
                 rsq_thr: float | None = None, *, out: str | None = None, extra: bool = False,
                 summary: bool = False, verbose: int = 0):
    lds = _module()
    bed, bim_path, fam_path = _bfile_paths(bfile)
    bim = pd.read_csv(bim_path, sep="\t", names=BIM_COLUMNS)
    if len(bim["CHR"].unique()) != 1:
        raise NLDSCParameterError("The current version of the program can only work with one chromosome in one file.")
    n_org = sum(1 for _ in open(fam_path))
    wind, metric = _window(ld_wind, wind_metric)
    maf = _unit_interval(maf_thr, 1.0, "Minor allele frequency")
    std = _unit_interval(std_thr, 1.0, "standard deviation threshold")
    rsq = _unit_interval(1.0 / len(bim) if rsq_thr is None else rsq_thr, 0.1, "r-squared threshold")
    positions = bim["CM" if metric == "cm" else "BP"].to_numpy(dtype=np.float64)
    params = lds.LDScoreParams(bfile=bed, n_snp=len(bim), n_org=n_org, ld_wind=wind, maf=maf, std_thr=std,
                               rsq_thr=rsq, positions=positions)
    ld = lds.calculate(params)
    if summary:
        d = pd.DataFrame({"L2": ld.l2, "L2D": ld.l2d, "MAF": ld.maf})
        print("L2/L2D/MAF Correlation matrix\n" + str(d.corr()))
        print(f"- Number of additive non-null LD: {d['L2'].count()}\n"
              f"- Number of non-additive non-null LD: {d['L2D'].count()}")
    out_df = make_output(bim, ld, extra=extra)
    if out:
        out_df.to_csv(out, sep="\t", index=False, float_format="%.5f")
        return None
    return out_df
