#!/bin/bash
# round 2, call ZF (1 GPU): end-to-end configs[4] call, 2 x 2-block body against the 160-column body, same box
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
AB_OUT=r02zf_cfg4_e2e_ab.json timeout 400 python scripts/cfg4_e2e_ab.py wide=NLDSC_X160=0 x160=NLDSC_X160=1 wide_again=NLDSC_X160=0 2>&1 | tail -4 | cut -c1-300
