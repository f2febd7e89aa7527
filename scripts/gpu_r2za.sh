#!/bin/bash
# round 2, call ZA (1 GPU): 160-column supertiles as the default from N = 6,000 -- whole GPU suite, bench + reference arm,
# and the same-box A/B around the threshold
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r02za_gpu_tests.txt 2>&1; echo "suite rc $?"; tail -4 gpurun_out/r02za_gpu_tests.txt | cut -c1-300
timeout 400 python bench.py --steps 20 --warmup 3 > gpurun_out/r02za_bench.json 2> gpurun_out/r02za_bench.err; echo "bench rc $?"
python -c "
import json; d=json.load(open('gpurun_out/r02za_bench.json')); r=d['roofline']
print('step', d['ms_per_step'], 'band', r['ms_band'], r['kernel'], 'frac', r['frac'], 'int8frac', r['frac_of_int8_peak'], 'exec', r['executed_frac'], r['executed_over_algorithmic'], 'e2e', d['e2e']['ms_per_step'], d.get('parity_check',{}).get('ok'), d['gpu_launches'])"
timeout 400 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02za_bench_reference.json 2> gpurun_out/r02za_bench_reference.err; echo "ref rc $?"
AB_OUT=r02za_x160_ab_n5000.json AB_N=5000 timeout 200 python scripts/ab_step.py wide=NLDSC_X160=0 x160=NLDSC_X160=1 2>&1 | tail -2 | cut -c1-200
AB_OUT=r02za_x160_ab_n7000.json AB_N=7000 timeout 200 python scripts/ab_step.py wide=NLDSC_X160=0 x160=NLDSC_X160=1 2>&1 | tail -2 | cut -c1-200
AB_OUT=r02za_x160_ab_n10000.json timeout 200 python scripts/ab_step.py wide=NLDSC_X160=0 x160=NLDSC_X160=1 2>&1 | tail -2 | cut -c1-200
