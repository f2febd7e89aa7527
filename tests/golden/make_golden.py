"""Generates tests/golden/*.npz: small synthetic bfiles plus the outputs of the
reference's OWN source compiled here (oracle/Makefile -> oracle/_ref/libref_ld64.so, the
FP64-promoted build, and libref_ld32.so, the stock float build).

The reference ships no tests, golden vectors or sample data, so these
fixtures are the pin: run in the build container (needs /root/reference), commit the .npz.

    python tests/golden/make_golden.py
"""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from nldsc_b200 import synth          # noqa: E402
from oracle import ref                # noqa: E402

# name, n_org, n_snp, missing, seed, density (SNPs per cM), window, metric, maf, std_thr
CASES = [
    ("n1000_m768_nomiss", 1000, 768, 0.0, 11, 340.0, 1.0, "cm", 1e-4, 1e-5),
    ("n1003_m600_miss", 1003, 600, 0.02, 3, 340.0, 1.0, "cm", 1e-4, 1e-5),     # n % 4 == 3
    ("n1001_m300_miss", 1001, 300, 0.05, 5, 150.0, 0.5, "cm", 1e-3, 1e-5),     # n % 4 == 1
    ("n1002_m300_kbp", 1002, 300, 0.01, 7, 340.0, 700.0, "kbp", 1e-4, 1e-4),   # n % 4 == 2, bp window
    ("n257_m1500_wide", 257, 1500, 0.03, 9, 340.0, 3.0, "cm", 0.01, 1e-5),     # window ~2000 SNPs > M
    ("n640_m200_tiny_window", 640, 200, 0.02, 13, 340.0, 0.001, "cm", 1e-4, 1e-5),
]


def main() -> None:
    here = os.path.dirname(os.path.abspath(__file__))
    for name, n_org, n_snp, missing, seed, density, wind, metric, maf, std_thr in CASES:
        spec = synth.SynthSpec(n_org=n_org, n_snp=n_snp, missing=missing, seed=seed, density=density)
        packed = synth.generate_packed(spec, workers=1)
        with tempfile.TemporaryDirectory() as d:
            info = synth.write_bfile(os.path.join(d, "g"), spec, packed)
            if metric == "cm":
                pos, w = info["cm"], wind
            else:
                pos, w = info["bp"].astype(np.float64), wind * 1000.0
            rsq_thr = 1.0 / n_snp
            out = {}
            for fl in (64, 32):
                r = ref.calculate(os.path.join(d, "g.bed"), n_snp, n_org, pos, w, maf, std_thr, rsq_thr, flavour=fl)
                for k, v in r.items():
                    out[f"ref{fl}_{k}"] = v
        np.savez_compressed(os.path.join(here, name + ".npz"), packed=packed, positions=pos,
                            n_org=n_org, ld_wind=w, maf=maf, std_thr=std_thr, rsq_thr=rsq_thr, **out)
        kept = int(np.sum(out["ref64_l2_ws"] >= 0))
        print(f"{name}: kept {kept}/{n_snp}, pairs {int(out['ref64_l2_ws'][out['ref64_l2_ws'] >= 0].sum())}")


if __name__ == "__main__":
    main()
