#!/bin/bash
# round 2, call J (1 GPU): one launch for both band bodies -- tests, A/B with and without missing genotypes, bench, sanitizer
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r02j_gpu_tests.txt 2>&1; echo "pytest rc $?" >> gpurun_out/r02j_gpu_tests.txt
tail -4 gpurun_out/r02j_gpu_tests.txt
AB_OUT=r02j_ab_missing_n10000.json timeout 300 python scripts/ab_step.py default mfree_off=NLDSC_MFREE=0 default_again 2>&1 | tail -3
AB_MISSING=0 AB_OUT=r02j_ab_mfree_n10000.json timeout 300 python scripts/ab_step.py default six_plane=NLDSC_MFREE=0 2>&1 | tail -3
AB_MISSING=0 AB_OUT=r02j_ab_mfree_n2504.json AB_N=2504 timeout 300 python scripts/ab_step.py default six_plane=NLDSC_MFREE=0 2>&1 | tail -3
AB_MISSING=0 AB_OUT=r02j_ab_mfree_cfg1.json AB_N=1000 AB_M=10000 timeout 300 python scripts/ab_step.py default six_plane=NLDSC_MFREE=0 2>&1 | tail -3
AB_MISSING=0 AB_OUT=r02j_ab_mfree_ukb.json AB_N=315599 AB_M=20000 AB_REPS=5 timeout 300 python scripts/ab_step.py default six_plane=NLDSC_MFREE=0 2>&1 | tail -3
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/r02j_bench.json 2> gpurun_out/r02j_bench.err; echo "bench rc $?"
python -c "
import json; d=json.load(open('gpurun_out/r02j_bench.json')); r=d['roofline']
print('step', d['ms_per_step'], 'band', r['ms_band'], 'frac', r['frac'], 'peak', r['peak'], 'exec', r['executed_frac'], 'e2e', d['e2e']['ms_per_step'], d['e2e'].get('bed_ms_warm'), d.get('parity_check',{}).get('ok'))"
tail -3 gpurun_out/r02j_bench.err
bash scripts/gpu_sanitize.sh
