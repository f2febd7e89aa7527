#!/bin/bash
# round 2, x160 call B (1 GPU): the 160-column supertiles against the default at other N
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
AB_OUT=x160_ab_n2504.json AB_N=2504 timeout 200 python scripts/ab_step.py default x160=NLDSC_X160=1 2>&1 | tail -2 | cut -c1-330
AB_OUT=x160_ab_n50000.json AB_N=50000 AB_M=40000 AB_REPS=10 timeout 200 python scripts/ab_step.py default x160=NLDSC_X160=1 2>&1 | tail -2 | cut -c1-330
AB_OUT=x160_ab_ukb.json AB_N=315599 AB_M=20000 AB_REPS=5 timeout 300 python scripts/ab_step.py default x160=NLDSC_X160=1 2>&1 | tail -2 | cut -c1-330
AB_OUT=x160_ab_n100k_10cm.json AB_N=100000 AB_M=40000 AB_WIND=10 AB_REPS=3 timeout 300 python scripts/ab_step.py default x160=NLDSC_X160=1 2>&1 | tail -2 | cut -c1-330
