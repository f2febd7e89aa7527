#!/bin/bash
cd "${GRAFT_REPO_ROOT:-.}"; mkdir -p gpurun_out
timeout 120 python - > gpurun_out/r02q_fp4_probe.txt 2>&1 <<'PY'
import numpy as np, json
from nldsc_b200 import capi
for n_mma in (1, 7, 6000, 150000):
    r = capi.debug_fp4_probe(n_mma, 256, 8)
    a = r["acc"]
    print("n_mma", n_mma, "ms", round(r["ms"], 3))
    print("  rank0 row0 cols0-15:", a[0, 0].tolist())
    print("  rank0 row77 cols:", a[0, 77, [0, 1, 8, 15]].tolist(), " rank1 row5 cols:", a[1, 5, [0, 1, 8, 15]].tolist())
    print("  uniform over rows:", bool((a == a[0, 0]).all()), " expected col0, col8, other:", n_mma, 4 * n_mma, 96 * n_mma)
for n_cols, pc in ((256, 8), (256, 4), (192, 8), (192, 16), (128, 8)):
    r = capi.debug_fp4_probe(200000, n_cols, pc, all_pairs=True)
    print("rate n_cols", n_cols, "per_commit", pc, "ms", round(r["ms"], 2), "TOP/s (2*M*N*K per MMA, K=64)", round(r["tops_equiv"], 1))
PY
echo "rc $?"; cat gpurun_out/r02q_fp4_probe.txt
