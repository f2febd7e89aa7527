#!/bin/bash
# round 2, x80 call C (1 GPU): the two-halves body where the 160-column body loses (small N) and at large N
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
AB_OUT=x80_ab_n2504_4d.json AB_N=2504 timeout 200 python scripts/ab_step.py wide=NLDSC_X160=0 x160=NLDSC_X160=1 x80=NLDSC_X160=1,NLDSC_X80=1 2>&1 | tail -3 | cut -c1-160
AB_OUT=x80_ab_n3500_4d.json AB_N=3500 timeout 200 python scripts/ab_step.py wide=NLDSC_X160=0 x160=NLDSC_X160=1 x80=NLDSC_X160=1,NLDSC_X80=1 2>&1 | tail -3 | cut -c1-160
AB_OUT=x80_ab_n50000_4d.json AB_N=50000 AB_M=40000 AB_REPS=10 timeout 200 python scripts/ab_step.py x160 x80=NLDSC_X80=1 2>&1 | tail -2 | cut -c1-160
