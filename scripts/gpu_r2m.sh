#!/bin/bash
# round 2, call M (1 GPU): final code -- full suite, default bench, ncu launch list (engine kernels only) and full captures of the
# band kernel on configs[1] (6-plane body) and on missing-free data (2-plane body)
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r02m_gpu_tests.txt 2>&1; echo "pytest rc $?" >> gpurun_out/r02m_gpu_tests.txt
tail -4 gpurun_out/r02m_gpu_tests.txt
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/r02m_bench.json 2> gpurun_out/r02m_bench.err; echo "bench rc $?"
python -c "
import json; d=json.load(open('gpurun_out/r02m_bench.json')); r=d['roofline']
print('step', d['ms_per_step'], 'band', r['ms_band'], 'frac', r['frac'], 'peak', r['peak'], 'exec', r['executed_frac'], 'e2e', d['e2e']['ms_per_step'], d['e2e'].get('bed_ms_warm'), d['e2e'].get('bed_ms_cold'), d.get('parity_check',{}).get('ok'))"
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02m_bench_reference.json 2> gpurun_out/r02m_bench_reference.err; echo "ref rc $?"; tail -c 600 gpurun_out/r02m_bench_reference.json
KN='regex:band_|unpack_|stats_kernel|finalize_kernel|windows_kernel|copy16_kernel|zero_rows|publish_rows|chunk_mfree|i8_peak'
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$KN" -c 600 --csv --log-file gpurun_out/r02m_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r02m_ncu_list.log 2>&1; echo "ncu list rc $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:band_tc_wide_kernel -s 20 -c 2 -o gpurun_out/r02m_band_full -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r02m_ncu_full.log 2>&1; echo "ncu full rc $?"
AB_MISSING=0 AB_REPS=3 timeout 900 ncu --set full --clock-control none --import-source on -k regex:band_tc_wide_kernel -s 8 -c 1 -o gpurun_out/r02m_band_mfree_full -f --target-processes all python scripts/ab_step.py default > gpurun_out/r02m_ncu_mfree.log 2>&1; echo "ncu mfree rc $?"
ls -la gpurun_out/*.ncu-rep
