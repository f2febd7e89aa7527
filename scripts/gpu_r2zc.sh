#!/bin/bash
# round 2, call ZC (1 GPU): final binary (160-column body from N = 4,096) -- whole GPU suite, bench + reference arm, ncu launch list
# and a full capture of band_tc_x160_kernel on configs[1], and the A/B below the threshold
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r02zc_gpu_tests.txt 2>&1; echo "suite rc $?"; tail -3 gpurun_out/r02zc_gpu_tests.txt | cut -c1-200
timeout 400 python bench.py --steps 20 --warmup 3 > gpurun_out/r02zc_bench.json 2> gpurun_out/r02zc_bench.err; echo "bench rc $?"
python -c "
import json; d=json.load(open('gpurun_out/r02zc_bench.json')); r=d['roofline']
print('step', d['ms_per_step'], 'band', r['ms_band'], r['kernel'], 'frac', r['frac'], 'int8frac', r['frac_of_int8_peak'], 'exec', r['executed_frac'], 'e2e', d['e2e']['ms_per_step'], d.get('parity_check',{}).get('ok'), d['gpu_launches'])"
timeout 400 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02zc_bench_reference.json 2> gpurun_out/r02zc_bench_reference.err; echo "ref rc $?"
KN='regex:band_|unpack_|stats_kernel|finalize_kernel|windows_kernel|copy16_kernel|zero_rows|publish_rows|chunk_mfree|i8_peak|fp4_probe'
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k "$KN" -c 600 --csv --log-file gpurun_out/r02zc_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r02zc_ncu_list.log 2>&1; echo "ncu list rc $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:band_tc_x160_kernel -s 20 -c 2 -o gpurun_out/r02zc_band_x160_full -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r02zc_ncu_full.log 2>&1; echo "ncu full rc $?"
AB_OUT=r02zc_x160_ab_n2504.json AB_N=2504 timeout 200 python scripts/ab_step.py wide=NLDSC_X160=0 x160=NLDSC_X160=1 2>&1 | tail -2 | cut -c1-120
AB_OUT=r02zc_x160_ab_n3500.json AB_N=3500 timeout 200 python scripts/ab_step.py wide=NLDSC_X160=0 x160=NLDSC_X160=1 2>&1 | tail -2 | cut -c1-120
ls -la gpurun_out/*.ncu-rep | tail -2
