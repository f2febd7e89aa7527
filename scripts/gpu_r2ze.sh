#!/bin/bash
# round 2, call ZE (1 GPU): the 160-column body on configs[4]-shaped work (dense map, 20,000-SNP windows)
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
AB_OUT=r02ze_x160_ab_cfg4.json AB_N=100000 AB_M=40000 AB_WIND=10 AB_DENSITY=1000 AB_REPS=3 timeout 400 python scripts/ab_step.py wide=NLDSC_X160=0 x160=NLDSC_X160=1 x160_r4=NLDSC_X160=1,NLDSC_RASTER=4 x160_r16=NLDSC_X160=1,NLDSC_RASTER=16 2>&1 | tail -4 | cut -c1-200
AB_OUT=r02ze_x160_ab_n10000_dense.json AB_N=10000 AB_M=60000 AB_WIND=10 AB_DENSITY=1000 AB_REPS=5 timeout 300 python scripts/ab_step.py wide=NLDSC_X160=0 x160=NLDSC_X160=1 2>&1 | tail -2 | cut -c1-200
