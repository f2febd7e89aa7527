# Soak: 600 fuzz seeds (60 per execution variant) against the oracle on one B200.
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
NLDSC_FUZZ_SEEDS=600 timeout 420 python -m pytest tests/test_gpu_fuzz.py -m gpu -q -p no:cacheprovider > gpurun_out/soak.log 2>&1; echo "soak rc=$?"
tail -5 gpurun_out/soak.log
