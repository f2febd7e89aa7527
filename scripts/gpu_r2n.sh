#!/bin/bash
# round 2, call N (1 GPU): paired commits in the missing-free body (8 MMAs per commit instead of 4) -- A/B
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
AB_MISSING=0 AB_OUT=r02n_ab_commit2_mfree_n10000.json timeout 300 python scripts/ab_step.py default commit2=NLDSC_COMMIT2=1 default_again 2>&1 | tail -3
AB_MISSING=0 AB_OUT=r02n_ab_commit2_mfree_ukb.json AB_N=315599 AB_M=20000 AB_REPS=5 timeout 300 python scripts/ab_step.py default commit2=NLDSC_COMMIT2=1 2>&1 | tail -2
AB_MISSING=0 AB_OUT=r02n_ab_commit2_mfree_n100k.json AB_N=100000 AB_M=40000 AB_DENSITY=1000 AB_WIND=10 AB_REPS=3 timeout 300 python scripts/ab_step.py default commit2=NLDSC_COMMIT2=1 six_plane=NLDSC_MFREE=0 2>&1 | tail -3
AB_MISSING=0 AB_OUT=r02n_ab_commit2_mfree_n2504.json AB_N=2504 timeout 300 python scripts/ab_step.py default commit2=NLDSC_COMMIT2=1 2>&1 | tail -2
