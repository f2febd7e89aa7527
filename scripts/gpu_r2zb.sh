#!/bin/bash
# round 2, call ZB (1 GPU): schedule knobs re-measured with the 160-column body (chunk rows, raster) and the end-to-end tail (taper)
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
AB_OUT=r02zb_x160_knobs_n10000.json timeout 300 python scripts/ab_step.py default c12k=NLDSC_CHUNK_ROWS=12288 c16k=NLDSC_CHUNK_ROWS=16384 c6k=NLDSC_CHUNK_ROWS=6144 r4=NLDSC_RASTER=4 r16=NLDSC_RASTER=16 default_again 2>&1 | tail -7 | cut -c1-150
for tm in default 2048 1024; do
  if [ $tm = default ]; then unset NLDSC_TAPER_MIN; else export NLDSC_TAPER_MIN=$tm; fi
  echo "taper_min $tm: $(timeout 120 python scripts/e2e_trace.py 2>&1 | grep 'e2e ms\|device:' | cut -c1-220 | tr '\n' ' ')"
done | tee gpurun_out/r02zb_e2e_taper.txt
