#!/bin/bash
# round 2, call X (1 GPU): fp4 accumulators -> double by one F2F each -- parity, then A/B against the float -> int -> double build
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fuzz.py -m gpu -x -q -k "golden or cfg1_full or wide or chunked or missing_free or unused or random_case or midsize" > gpurun_out/r02x_quick.txt 2>&1; rc=$?; echo "quick rc $rc"; tail -4 gpurun_out/r02x_quick.txt | cut -c1-300
L=$PWD/nldsc_b200/libnldsc_ld_i2d.so
AB_OUT=r02x_ab_f2f_n10000.json timeout 300 python scripts/ab_step.py default i2d=NLDSC_LIB=$L default_again i2d_again=NLDSC_LIB=$L 2>&1 | tail -4 | cut -c1-200
AB_OUT=r02x_ab_f2f_n2504.json AB_N=2504 timeout 300 python scripts/ab_step.py default i2d=NLDSC_LIB=$L 2>&1 | tail -2 | cut -c1-200
AB_MISSING=0 AB_OUT=r02x_ab_f2f_mfree_n10000.json timeout 300 python scripts/ab_step.py default i2d=NLDSC_LIB=$L 2>&1 | tail -2 | cut -c1-200
AB_OUT=r02x_ab_f2f_ukb.json AB_N=315599 AB_M=20000 AB_REPS=5 timeout 300 python scripts/ab_step.py default i2d=NLDSC_LIB=$L 2>&1 | tail -2 | cut -c1-200
python scripts/e2e_trace.py 2>&1 | grep "e2e ms"
