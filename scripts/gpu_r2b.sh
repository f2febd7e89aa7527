#!/bin/bash
# round 2, call B (2 GPUs): full GPU suite on a multi-GPU box, bench --gpus 2 (strong record), genome driver phases,
# a small cut of the UKB genome script, the int8 peak sweep, and the wide MMA-only probe mode under a timeout
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > gpurun_out/r02b_box.txt; nproc >> gpurun_out/r02b_box.txt; free -g >> gpurun_out/r02b_box.txt; df -h /tmp >> gpurun_out/r02b_box.txt
timeout 1200 python -m pytest tests -m gpu -x -q --durations=8 > gpurun_out/r02b_gpu_tests.txt 2>&1; echo "pytest rc $?" >> gpurun_out/r02b_gpu_tests.txt
tail -14 gpurun_out/r02b_gpu_tests.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/r02b_bench_n2.json 2> gpurun_out/r02b_bench_n2.err; echo "bench n2 rc $?"
tail -c 2500 gpurun_out/r02b_bench_n2.json; tail -5 gpurun_out/r02b_bench_n2.err
timeout 300 python - > gpurun_out/r02b_i8_peak.json 2> gpurun_out/r02b_i8_peak.err <<'PY'
import json
from nldsc_b200 import capi
out = []
for n_cols, pc, depth, iters in [(192, 8, 4, 100000), (192, 8, 8, 100000), (192, 16, 4, 50000), (192, 4, 8, 200000), (256, 8, 4, 75000), (256, 16, 4, 40000), (128, 8, 4, 150000), (128, 16, 8, 75000), (96, 16, 8, 75000), (64, 16, 8, 75000)]:
    out.append(capi.debug_i8_peak(n_cols, pc, iters, reps=3, depth=depth))
print(json.dumps(out))
PY
cat gpurun_out/r02b_i8_peak.json; tail -3 gpurun_out/r02b_i8_peak.err
for g in 1 2; do NLDSC_TRACE=1 timeout 600 python scripts/cfg3_genome.py $g 50000 > gpurun_out/r02b_cfg3_${g}gpu.log 2> gpurun_out/r02b_cfg3_${g}gpu.err; echo "cfg3 $g rc $?"; tail -c 1500 gpurun_out/r02b_cfg3_${g}gpu.log; grep "genome:" gpurun_out/r02b_cfg3_${g}gpu.err | tail -2; done
timeout 900 python scripts/ukb_genome.py 2 80000 > gpurun_out/r02b_ukb_small.log 2> gpurun_out/r02b_ukb_small.err; echo "ukb rc $?"; tail -c 2500 gpurun_out/r02b_ukb_small.log; tail -5 gpurun_out/r02b_ukb_small.err
timeout 120 python -c "
from nldsc_b200 import capi
print('mode5', capi.debug_band_probe(5, 10000, 2000, 3))
" > gpurun_out/r02b_mode5.txt 2>&1; echo "mode5 rc $?"; tail -3 gpurun_out/r02b_mode5.txt
