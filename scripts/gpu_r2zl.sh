#!/bin/bash
# round 2, call ZL (1 GPU): final binary (PRMT lane selects in the 160-column body only) -- the tests that reach that body, the
# compiled-reference fuzz cases and the allocation-free entry point, then the bench
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "x160 or golden or cfg1_full or cfg4 or cfg5 or random_case or midsize or unused or sharded or zz or static" > gpurun_out/r02zl_gpu_tests_subset.txt 2>&1; echo "subset rc $?"; tail -2 gpurun_out/r02zl_gpu_tests_subset.txt | cut -c1-200
timeout 300 python bench.py --steps 20 --warmup 3 > gpurun_out/r02zl_bench.json 2> gpurun_out/r02zl_bench.err; echo "bench rc $?"
python -c "
import json; d=json.load(open('gpurun_out/r02zl_bench.json')); r=d['roofline']
print('step', d['ms_per_step'], 'band', r['ms_band'], r['kernel'], 'frac', r['frac'], 'int8frac', r['frac_of_int8_peak'], 'e2e', d['e2e']['ms_per_step'], d.get('parity_check',{}).get('ok'), d['gpu_launches'])"
