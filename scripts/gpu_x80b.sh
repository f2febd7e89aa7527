#!/bin/bash
# round 2, x80 call B (1 GPU): A and B staged by one 4-D TMA box each (2 bulk-tensor operations per stage instead of 9) --
# bitwise checks of both bodies (including the wrapping-ring fallback), then the same-box A/B
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
X160_CHECK_ENV="NLDSC_X160=1" timeout 200 python scripts/x160_check.py > gpurun_out/x160_check_4d.txt 2>&1; echo "x160 check rc $?"; grep -c OK gpurun_out/x160_check_4d.txt; grep -v " OK " gpurun_out/x160_check_4d.txt | tail -5 | cut -c1-300
X160_CHECK_ENV="NLDSC_X160=1,NLDSC_X80=1" timeout 200 python scripts/x160_check.py > gpurun_out/x80_check_4d.txt 2>&1; echo "x80 check rc $?"; grep -c OK gpurun_out/x80_check_4d.txt; grep -v " OK " gpurun_out/x80_check_4d.txt | tail -5 | cut -c1-300
AB_OUT=x80_ab_n10000_4d.json timeout 200 python scripts/ab_step.py x160 x80=NLDSC_X80=1 wide=NLDSC_X160=0 2>&1 | tail -3 | cut -c1-250
AB_OUT=x80_ab_n5000_4d.json AB_N=5000 timeout 200 python scripts/ab_step.py x160 x80=NLDSC_X80=1 2>&1 | tail -2 | cut -c1-250
