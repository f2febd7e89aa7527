#!/bin/bash
# round 2, call F (1 GPU): fused launch after the lazy-loading fix -- quick fuzz subset first, then A/B, full tests, bench
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_fuzz.py -m gpu -x -q -k "random_case" > gpurun_out/r02f_fuzz.txt 2>&1; rc=$?; echo "fuzz rc $rc"; tail -3 gpurun_out/r02f_fuzz.txt
if [ $rc -ne 0 ]; then export NLDSC_FUSED=0; echo "FUSED DISABLED FOR THE REST"; fi
AB_OUT=r02f_ab_step.json timeout 600 python scripts/ab_step.py default fused0=NLDSC_FUSED=0 default_noskip=NLDSC_NOSKIP=1 default_again fused0_again=NLDSC_FUSED=0 2>&1 | tail -6
AB_OUT=r02f_ab_step_n2504.json AB_N=2504 timeout 300 python scripts/ab_step.py default fused0=NLDSC_FUSED=0 2>&1 | tail -3
AB_OUT=r02f_ab_step_n1000.json AB_N=1000 AB_M=10000 timeout 300 python scripts/ab_step.py default fused0=NLDSC_FUSED=0 2>&1 | tail -3
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r02f_gpu_tests.txt 2>&1; echo "pytest rc $?" >> gpurun_out/r02f_gpu_tests.txt
tail -4 gpurun_out/r02f_gpu_tests.txt
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/r02f_bench.json 2> gpurun_out/r02f_bench.err; echo "bench rc $?"
python -c "
import json; d=json.load(open('gpurun_out/r02f_bench.json')); r=d['roofline']
print('step', d['ms_per_step'], 'band', r['ms_band'], 'frac', r['frac'], 'peak', r['peak'], 'exec', r['executed_frac'], 'e2e', d['e2e']['ms_per_step'], d['e2e'].get('bed_ms_warm'), d.get('parity_check',{}).get('ok'))"
tail -3 gpurun_out/r02f_bench.err
