#!/bin/bash
# round 2, call T (1 GPU): accumulator-0 run-ahead in the wide body -- parity, then A/B
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fuzz.py -m gpu -x -q -k "golden or cfg1_full or wide or chunked or ragged or random_case or midsize" > gpurun_out/r02t_quick.txt 2>&1; rc=$?; echo "quick rc $rc"; tail -4 gpurun_out/r02t_quick.txt | cut -c1-300
AB_OUT=r02t_ab_runahead_n10000.json timeout 300 python scripts/ab_step.py default off=NLDSC_RUNAHEAD=0 default_again off_again=NLDSC_RUNAHEAD=0 2>&1 | tail -4 | cut -c1-200
AB_OUT=r02t_ab_runahead_n2504.json AB_N=2504 timeout 300 python scripts/ab_step.py default off=NLDSC_RUNAHEAD=0 2>&1 | tail -2 | cut -c1-200
AB_OUT=r02t_ab_runahead_ukb.json AB_N=315599 AB_M=20000 AB_REPS=5 timeout 300 python scripts/ab_step.py default off=NLDSC_RUNAHEAD=0 2>&1 | tail -2 | cut -c1-200
AB_OUT=r02t_ab_runahead_int8.json timeout 300 python scripts/ab_step.py int8=NLDSC_FP4=0 int8_off=NLDSC_FP4=0,NLDSC_RUNAHEAD=0 2>&1 | tail -2 | cut -c1-200
