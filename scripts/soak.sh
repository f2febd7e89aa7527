# Soak on one B200: 600 small fuzz seeds (60 per execution variant) and 30 multi-chunk cases against the oracle.
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}"; mkdir -p gpurun_out
NLDSC_FUZZ_SEEDS=${NLDSC_FUZZ_SEEDS:-600} NLDSC_FUZZ_MID=${NLDSC_FUZZ_MID:-30} timeout 900 python -m pytest tests/test_gpu_fuzz.py -m gpu -q -p no:cacheprovider > gpurun_out/soak.log 2>&1; echo "soak rc=$?"
grep -n "^E  .*Error\|^FAILED" gpurun_out/soak.log | head -20; tail -3 gpurun_out/soak.log
