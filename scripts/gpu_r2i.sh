#!/bin/bash
# round 2, call I (1 GPU): missing-free 2-plane kernel -- targeted tests first, then A/B on missing-free workloads, full suite
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "missing_free or cfg1_full or golden or degenerate" > gpurun_out/r02i_mfree_tests.txt 2>&1; rc=$?; echo "mfree tests rc $rc"; tail -15 gpurun_out/r02i_mfree_tests.txt
# A/B needs missing-free data: ab_step generates with 2% missing; AB_MISSING=0 switches it off
AB_MISSING=0 AB_OUT=r02i_ab_mfree_n2504.json AB_N=2504 timeout 300 python scripts/ab_step.py default six_plane=NLDSC_MFREE=0 2>&1 | tail -3
AB_MISSING=0 AB_OUT=r02i_ab_mfree_cfg1.json AB_N=1000 AB_M=10000 timeout 300 python scripts/ab_step.py default six_plane=NLDSC_MFREE=0 2>&1 | tail -3
AB_MISSING=0 AB_OUT=r02i_ab_mfree_n10000.json timeout 300 python scripts/ab_step.py default six_plane=NLDSC_MFREE=0 2>&1 | tail -3
AB_MISSING=0 AB_OUT=r02i_ab_mfree_ukb.json AB_N=315599 AB_M=20000 AB_REPS=5 timeout 300 python scripts/ab_step.py default six_plane=NLDSC_MFREE=0 2>&1 | tail -3
AB_OUT=r02i_ab_missing_n10000.json timeout 300 python scripts/ab_step.py default mfree_off=NLDSC_MFREE=0 2>&1 | tail -3
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r02i_gpu_tests.txt 2>&1; echo "pytest rc $?" >> gpurun_out/r02i_gpu_tests.txt
tail -4 gpurun_out/r02i_gpu_tests.txt
