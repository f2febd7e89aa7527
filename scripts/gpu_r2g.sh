#!/bin/bash
# round 2, call G (1 GPU): column constants staged in shared memory (wide kernel epilogue) -- A/B on the same box
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_fuzz.py tests/test_gpu_parity.py -m gpu -x -q -k "random_case or golden or wide or chunked or cfg1" > gpurun_out/r02g_quick.txt 2>&1; echo "quick rc $?"; tail -3 gpurun_out/r02g_quick.txt
L=$PWD/nldsc_b200
AB_OUT=r02g_ab_step.json timeout 600 python scripts/ab_step.py default nostage=NLDSC_LIB=$L/libnldsc_ld_nostage.so v1epi=NLDSC_LIB=$L/libnldsc_ld_v1epi.so v1epi_nostage=NLDSC_LIB=$L/libnldsc_ld_v1epi_nostage.so default_again 2>&1 | tail -6
AB_OUT=r02g_ab_step_n2504.json AB_N=2504 timeout 300 python scripts/ab_step.py default nostage=NLDSC_LIB=$L/libnldsc_ld_nostage.so v1epi=NLDSC_LIB=$L/libnldsc_ld_v1epi.so 2>&1 | tail -4
AB_OUT=r02g_ab_step_n1000.json AB_N=1000 AB_M=10000 timeout 300 python scripts/ab_step.py default nostage=NLDSC_LIB=$L/libnldsc_ld_nostage.so v1epi=NLDSC_LIB=$L/libnldsc_ld_v1epi.so 2>&1 | tail -4
AB_OUT=r02g_ab_step_ukb.json AB_N=315599 AB_M=20000 AB_REPS=5 timeout 300 python scripts/ab_step.py default nostage=NLDSC_LIB=$L/libnldsc_ld_nostage.so 2>&1 | tail -3
timeout 200 python - > gpurun_out/r02g_i8_peak.json 2> gpurun_out/r02g_i8_peak.err <<'PY'
import json
from nldsc_b200 import capi
out = []
for n_cols, pc, depth, iters in [(192, 8, 4, 100000), (192, 8, 4 | 0x100, 100000), (192, 4, 4 | 0x100, 200000), (192, 16, 4, 50000), (192, 16, 4 | 0x100, 50000), (256, 8, 4 | 0x100, 75000)]:
    r = capi.debug_i8_peak(n_cols, pc, iters, reps=3, depth=depth)
    r["commit"] = "leader-only" if depth & 0x100 else "multicast"
    out.append(r)
print(json.dumps(out))
PY
cat gpurun_out/r02g_i8_peak.json; tail -2 gpurun_out/r02g_i8_peak.err
