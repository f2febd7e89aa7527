#!/bin/bash
# round 2, x160 call A (1 GPU): 160-column supertiles -- bitwise check against the default, then the same-box A/B
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 240 python scripts/x160_check.py > gpurun_out/x160_check.txt 2>&1; echo "check rc $?"; tail -14 gpurun_out/x160_check.txt | cut -c1-260
AB_OUT=x160_ab_n10000.json timeout 200 python scripts/ab_step.py default x160=NLDSC_X160=1 default_again x160_again=NLDSC_X160=1 2>&1 | tail -4 | cut -c1-330
