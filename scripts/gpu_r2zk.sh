#!/bin/bash
# round 2, call ZK (1 GPU): one-instruction lane selects (PRMT) in the 6-plane epilogues -- A/B against the previous build, then the
# whole GPU suite and the bench on the final binary
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
P=$PWD/nldsc_b200/libnldsc_ld_prev.so
AB_OUT=r02zk_ab_prmt_n10000.json timeout 200 python scripts/ab_step.py default prev=NLDSC_LIB=$P default_again 2>&1 | tail -3 | cut -c1-200
AB_OUT=r02zk_ab_prmt_n2504.json AB_N=2504 timeout 200 python scripts/ab_step.py default prev=NLDSC_LIB=$P 2>&1 | tail -2 | cut -c1-200
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r02zk_gpu_tests.txt 2>&1; echo "suite rc $?"; tail -2 gpurun_out/r02zk_gpu_tests.txt | cut -c1-200
timeout 400 python bench.py --steps 20 --warmup 3 > gpurun_out/r02zk_bench.json 2> gpurun_out/r02zk_bench.err; echo "bench rc $?"
python -c "
import json; d=json.load(open('gpurun_out/r02zk_bench.json')); r=d['roofline']
print('step', d['ms_per_step'], 'band', r['ms_band'], r['kernel'], 'frac', r['frac'], 'int8frac', r['frac_of_int8_peak'], 'e2e', d['e2e']['ms_per_step'], d.get('parity_check',{}).get('ok'), d['gpu_launches'])"
