#!/bin/bash
# round 2, call Y (1 GPU): A/B of the band kernel's register cap.  Build the variants first, here:
#   python -c "from nldsc_b200 import build; build.build_variant('r144', ['-DNLDSC_WIDE_MAXNREG=144']); build.build_variant('r160', ['-DNLDSC_WIDE_MAXNREG=160'])"
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
A=$PWD/nldsc_b200/libnldsc_ld_r144.so
B=$PWD/nldsc_b200/libnldsc_ld_r160.so
AB_OUT=r02y_ab_regs_n10000.json timeout 300 python scripts/ab_step.py default r144=NLDSC_LIB=$A r160=NLDSC_LIB=$B default_again 2>&1 | tail -4 | cut -c1-200
AB_OUT=r02y_ab_regs_n2504.json AB_N=2504 timeout 300 python scripts/ab_step.py default r144=NLDSC_LIB=$A r160=NLDSC_LIB=$B 2>&1 | tail -3 | cut -c1-200
AB_OUT=r02y_ab_regs_ukb.json AB_N=315599 AB_M=20000 AB_REPS=5 timeout 300 python scripts/ab_step.py default r144=NLDSC_LIB=$A r160=NLDSC_LIB=$B 2>&1 | tail -3 | cut -c1-200
