#!/bin/bash
# round 2, x80 call A (1 GPU): two-halves body -- bitwise check against the 2 x 2-block body, then the same-box A/B
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
X160_CHECK_ENV="NLDSC_X160=1,NLDSC_X80=1" timeout 240 python scripts/x160_check.py > gpurun_out/x80_check.txt 2>&1; echo "check rc $?"; tail -12 gpurun_out/x80_check.txt | cut -c1-200
AB_OUT=x80_ab_n10000.json timeout 200 python scripts/ab_step.py x160 x80=NLDSC_X80=1 wide=NLDSC_X160=0 x80_again=NLDSC_X80=1 2>&1 | tail -4 | cut -c1-250
AB_OUT=x80_ab_n5000.json AB_N=5000 timeout 200 python scripts/ab_step.py x160 x80=NLDSC_X80=1 2>&1 | tail -2 | cut -c1-250
