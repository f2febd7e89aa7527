#!/bin/bash
# round 2, call ZI (1 GPU): upload piece size against the end-to-end call on configs[1]
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
for pb in default 4194304 16777216 33554432 67108864; do
  if [ $pb = default ]; then unset NLDSC_PIECE_BYTES; else export NLDSC_PIECE_BYTES=$pb; fi
  echo "piece_bytes $pb: $(timeout 120 python scripts/e2e_trace.py 2>&1 | grep 'e2e ms\|device:' | cut -c1-200 | tr '\n' ' ')"
done | tee gpurun_out/r02zi_e2e_piece.txt
