#!/bin/bash
# compute-sanitizer over the golden cases: memcheck, racecheck, synccheck -- each under its own timeout
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
python scripts/sanitize.py > gpurun_out/r02_sanitize_plain.txt 2>&1; echo "plain rc $?"
for tool in memcheck racecheck synccheck; do
  timeout 600 compute-sanitizer --tool $tool --print-limit 20 python scripts/sanitize.py > gpurun_out/r02_sanitize_$tool.txt 2>&1
  echo "$tool rc $?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|error|hazard" gpurun_out/r02_sanitize_$tool.txt | tail -5
done
