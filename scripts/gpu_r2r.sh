#!/bin/bash
# round 2, call R (1 GPU): fp4 band bodies -- unpack + targeted parity first, then A/B against the int8 bodies, full suite, bench
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 500 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "fp4_unpack or golden or cfg1_full or missing_free or wide or chunked or ragged or degenerate or unused" > gpurun_out/r02r_quick.txt 2>&1; rc=$?; echo "quick rc $rc"; tail -12 gpurun_out/r02r_quick.txt | cut -c1-300
AB_OUT=r02r_ab_fp4_n10000.json timeout 300 python scripts/ab_step.py default int8=NLDSC_FP4=0 default_again 2>&1 | tail -3 | cut -c1-330
AB_MISSING=0 AB_OUT=r02r_ab_fp4_mfree_n10000.json timeout 300 python scripts/ab_step.py default int8=NLDSC_FP4=0 2>&1 | tail -2 | cut -c1-330
AB_OUT=r02r_ab_fp4_n2504.json AB_N=2504 timeout 300 python scripts/ab_step.py default int8=NLDSC_FP4=0 2>&1 | tail -2 | cut -c1-330
AB_OUT=r02r_ab_fp4_ukb.json AB_N=315599 AB_M=20000 AB_REPS=5 timeout 300 python scripts/ab_step.py default int8=NLDSC_FP4=0 2>&1 | tail -2 | cut -c1-330
AB_OUT=r02r_ab_fp4_n100k.json AB_N=100000 AB_M=40000 AB_DENSITY=1000 AB_WIND=10 AB_REPS=3 timeout 300 python scripts/ab_step.py default int8=NLDSC_FP4=0 2>&1 | tail -2 | cut -c1-330
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r02r_gpu_tests.txt 2>&1; echo "pytest rc $?" >> gpurun_out/r02r_gpu_tests.txt
tail -5 gpurun_out/r02r_gpu_tests.txt | cut -c1-300
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/r02r_bench.json 2> gpurun_out/r02r_bench.err; echo "bench rc $?"
python -c "
import json; d=json.load(open('gpurun_out/r02r_bench.json')); r=d['roofline']
print('step', d['ms_per_step'], 'band', r['ms_band'], 'frac', r['frac'], 'peak', r['peak'], 'frac_int8', r['frac_of_int8_peak'], 'exec', r['executed_frac'], 'e2e', d['e2e']['ms_per_step'], d.get('parity_check',{}).get('ok'))"
tail -3 gpurun_out/r02r_bench.err
