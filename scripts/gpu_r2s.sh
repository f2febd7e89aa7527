#!/bin/bash
# round 2, call S (1 GPU): re-tune the schedule knobs now that the band kernel runs on the fp4 pipe
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
AB_OUT=r02s_ab_knobs_n10000.json timeout 400 python scripts/ab_step.py default c4k=NLDSC_CHUNK_ROWS=4096 c16k=NLDSC_CHUNK_ROWS=16384 r4=NLDSC_RASTER=4 r16=NLDSC_RASTER=16 bs1=NLDSC_BAND_STREAMS=1 noskip=NLDSC_NOSKIP=1 fused=NLDSC_FUSED=1 2>&1 | tail -8 | cut -c1-200
AB_OUT=r02s_ab_knobs_ukb.json AB_N=315599 AB_M=20000 AB_REPS=5 timeout 300 python scripts/ab_step.py default overlap=NLDSC_WIDE_OVERLAP=1 c4k=NLDSC_CHUNK_ROWS=4096 2>&1 | tail -3 | cut -c1-200
AB_OUT=r02s_ab_knobs_n100k.json AB_N=100000 AB_M=40000 AB_DENSITY=1000 AB_WIND=10 AB_REPS=3 timeout 300 python scripts/ab_step.py default overlap=NLDSC_WIDE_OVERLAP=1 r4=NLDSC_RASTER=4 r16=NLDSC_RASTER=16 2>&1 | tail -4 | cut -c1-200
AB_OUT=r02s_ab_knobs_n2504.json AB_N=2504 timeout 300 python scripts/ab_step.py default c16k=NLDSC_CHUNK_ROWS=16384 c64k=NLDSC_CHUNK_ROWS=65536 2>&1 | tail -3 | cut -c1-200
timeout 100 python -c "
from nldsc_b200 import capi
import json
out={}
for name,mode,n,t in [('n10000_wide_real_int8',6,10000,74*80),('n10000_wide_mma_only_int8',5,10000,74*80)]:
    out[name]=capi.debug_band_probe(mode,n,t,3)
print(json.dumps(out))
" 2>&1 | tail -2
