#!/bin/bash
# round 2, call C (1 GPU): new epilogue math + group skipping; paired-commit A/B; ncu launch list + full capture of the band kernel
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --durations=5 > gpurun_out/r02c_gpu_tests.txt 2>&1; echo "pytest rc $?" >> gpurun_out/r02c_gpu_tests.txt
tail -9 gpurun_out/r02c_gpu_tests.txt
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/r02c_bench.json 2> gpurun_out/r02c_bench.err; echo "bench rc $?"
python -c "
import json; d=json.load(open('gpurun_out/r02c_bench.json')); r=d['roofline']
print('step', d['ms_per_step'], 'band', r['ms_band'], 'frac', r['frac'], 'peak', r['peak'], 'exec', r['executed_frac'], 'e2e', d['e2e']['ms_per_step'], d.get('parity_check',{}).get('ok'))"
NLDSC_COMMIT2=1 timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r02c_bench_commit2.json 2> gpurun_out/r02c_bench_commit2.err; echo "bench commit2 rc $?"
python -c "
import json; d=json.load(open('gpurun_out/r02c_bench_commit2.json')); r=d['roofline']
print('commit2: step', d['ms_per_step'], 'band', r['ms_band'], 'frac', r['frac'], 'e2e', d['e2e']['ms_per_step'])"
timeout 200 python -c "
from nldsc_b200 import capi
import json
out={}
for name,mode,n,t in [('n10000_wide_real',6,10000,6000),('n10000_wide_mma_only',5,10000,6000),('n100000_wide_real',6,100000,1500),('n100000_wide_mma_only',5,100000,1500),('n10000_narrow_mma_only',1,10000,6000)]:
    out[name]=capi.debug_band_probe(mode,n,t,3)
print(json.dumps(out))
" > gpurun_out/r02c_band_probe.json 2> gpurun_out/r02c_band_probe.err; echo "probe rc $?"; cat gpurun_out/r02c_band_probe.json; tail -2 gpurun_out/r02c_band_probe.err
# launch list (serialised, cold) of one bench run, then one full capture of the band kernel
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02c_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r02c_ncu_list.log 2>&1; echo "ncu list rc $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:band_tc_wide_kernel -s 20 -c 2 -o gpurun_out/r02c_band_full -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r02c_ncu_full.log 2>&1; echo "ncu full rc $?"
ls -la gpurun_out/*.ncu-rep 2>/dev/null
