#!/bin/bash
# round 2, call K (1 GPU): chunk-size A/B at small N, the sanitize targets without the sanitizer (closed on this pool)
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
python scripts/sanitize.py > gpurun_out/r02k_sanitize_targets_plain.txt 2>&1; echo "plain rc $?"; tail -2 gpurun_out/r02k_sanitize_targets_plain.txt
AB_MISSING=0 AB_OUT=r02k_ab_chunks_mfree_n2504.json AB_N=2504 timeout 300 python scripts/ab_step.py default c16k=NLDSC_CHUNK_ROWS=16384 c32k=NLDSC_CHUNK_ROWS=32768 c64k=NLDSC_CHUNK_ROWS=65536 2>&1 | tail -4
AB_OUT=r02k_ab_chunks_n2504.json AB_N=2504 timeout 300 python scripts/ab_step.py default c16k=NLDSC_CHUNK_ROWS=16384 c32k=NLDSC_CHUNK_ROWS=32768 2>&1 | tail -3
AB_OUT=r02k_ab_chunks_n10000.json timeout 300 python scripts/ab_step.py default c16k=NLDSC_CHUNK_ROWS=16384 c4k=NLDSC_CHUNK_ROWS=4096 2>&1 | tail -3
AB_MISSING=0 AB_OUT=r02k_ab_chunks_mfree_n10000.json timeout 300 python scripts/ab_step.py default c16k=NLDSC_CHUNK_ROWS=16384 c32k=NLDSC_CHUNK_ROWS=32768 2>&1 | tail -3
AB_MISSING=0 AB_OUT=r02k_ab_chunks_mfree_cfg1.json AB_N=1000 AB_M=10000 timeout 300 python scripts/ab_step.py default c16k=NLDSC_CHUNK_ROWS=16384 2>&1 | tail -2
