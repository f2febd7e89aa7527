#!/bin/bash
# round 2, call E (1 GPU): fused single band launch -- tests, then same-box A/B of the step
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r02e_gpu_tests.txt 2>&1; echo "pytest rc $?" >> gpurun_out/r02e_gpu_tests.txt
tail -4 gpurun_out/r02e_gpu_tests.txt
V1=$PWD/nldsc_b200/libnldsc_ld_v1epi.so
AB_OUT=r02e_ab_step.json timeout 600 python scripts/ab_step.py default fused0=NLDSC_FUSED=0 noskip=NLDSC_NOSKIP=1 fused0_noskip=NLDSC_FUSED=0,NLDSC_NOSKIP=1 \
   v1epi=NLDSC_LIB=$V1 v1epi_fused0=NLDSC_LIB=$V1,NLDSC_FUSED=0 v1epi_fused0_noskip=NLDSC_LIB=$V1,NLDSC_FUSED=0,NLDSC_NOSKIP=1 default_again 2>&1 | tail -12
AB_OUT=r02e_ab_step_n2504.json AB_N=2504 timeout 300 python scripts/ab_step.py default fused0=NLDSC_FUSED=0 2>&1 | tail -3
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/r02e_bench.json 2> gpurun_out/r02e_bench.err; echo "bench rc $?"
python -c "
import json; d=json.load(open('gpurun_out/r02e_bench.json')); r=d['roofline']
print('step', d['ms_per_step'], 'band', r['ms_band'], 'frac', r['frac'], 'peak', r['peak'], 'exec', r['executed_frac'], 'e2e', d['e2e']['ms_per_step'], d['e2e'].get('bed_ms_warm'), d.get('parity_check',{}).get('ok'))"
tail -3 gpurun_out/r02e_bench.err
