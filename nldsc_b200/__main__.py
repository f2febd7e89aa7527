"""`python -m nldsc_b200 ld ...` -- the reference's `ld` command line (nldsc/__main__.py:35-97):
same flags, same exactly-one-window rule, same output file."""
from __future__ import annotations

import argparse
import sys

from .ldscore import estimate_lds, estimate_lds_genome


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="nldsc_b200")
    sub = ap.add_subparsers(dest="cmd", required=True)
    ld = sub.add_parser("ld", help="estimate additive and dominance LD scores")
    ld.add_argument("--bfile", required=True)
    ld.add_argument("-o", "--out")
    ld.add_argument("-kb", "--ld-wind-kb", dest="ld_wind_kb")
    ld.add_argument("-cm", "--ld-wind-cm", dest="ld_wind_cm")
    ld.add_argument("-maf", "--maf-thr", dest="maf_thr", required=True)   # effectively mandatory
    ld.add_argument("-std", "--std-thr", dest="std_thr", default="1e-4")
    ld.add_argument("-rsq", "--rsq-thr", dest="rsq_thr", default=None)
    ld.add_argument("--extra", action="store_true")
    ld.add_argument("--display", action="store_true")
    ld.add_argument("--m-files", action="store_true", help="also write <out stem>.M and .M_5_50 (not in the reference)")
    # not in the reference: every chromosome of a genome in one process, spread over the GPUs
    gn = sub.add_parser("ld-genome", help="LD scores of all chromosome bfiles matching prefix@suffix")
    gn.add_argument("--bfile", required=True, help="bfile pattern, @ = chromosome number")
    gn.add_argument("--out-dir", required=True)
    gn.add_argument("-kb", "--ld-wind-kb", dest="ld_wind_kb")
    gn.add_argument("-cm", "--ld-wind-cm", dest="ld_wind_cm")
    gn.add_argument("-maf", "--maf-thr", dest="maf_thr", required=True)
    gn.add_argument("-std", "--std-thr", dest="std_thr", default="1e-4")
    gn.add_argument("-rsq", "--rsq-thr", dest="rsq_thr", default=None)
    gn.add_argument("--gpus", type=int, default=None)
    a = ap.parse_args(argv)
    if (a.ld_wind_kb is None) == (a.ld_wind_cm is None):
        print("Must specify exactly one --ld-wind option", file=sys.stderr)
        return 2
    wind, metric = (a.ld_wind_kb, "kbp") if a.ld_wind_kb is not None else (a.ld_wind_cm, "cm")
    try:
        if a.cmd == "ld-genome":
            print(estimate_lds_genome(a.bfile, wind, metric, a.maf_thr, a.std_thr, a.rsq_thr, out_dir=a.out_dir,
                                      gpus=a.gpus).to_string(index=False))
            return 0
        df = estimate_lds(a.bfile, wind, metric, a.maf_thr, a.std_thr, a.rsq_thr, out=a.out, extra=a.extra,
                          summary=True, m_files=a.m_files)
    except Exception as e:            # reference handle_exception: log + SystemExit (__main__.py:17-26)
        print(f"error: {e}", file=sys.stderr)
        return 1
    if df is not None and a.display:
        print(df.to_string())
    return 0


if __name__ == "__main__":
    sys.exit(main())
