#!/bin/bash
# round 2, call A (1 GPU): GPU tests, the default bench, the int8 peak sweep
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > gpurun_out/r02a_box.txt; nproc >> gpurun_out/r02a_box.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r02a_gpu_tests.txt 2>&1; echo "pytest rc $?" >> gpurun_out/r02a_gpu_tests.txt
tail -3 gpurun_out/r02a_gpu_tests.txt
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/r02a_bench.json 2> gpurun_out/r02a_bench.err; echo "bench rc $?"
tail -c 3000 gpurun_out/r02a_bench.json
timeout 300 python - > gpurun_out/r02a_i8_peak.json 2> gpurun_out/r02a_i8_peak.err <<'PY'
import json
from nldsc_b200 import capi
out = []
for n_cols, pc, iters in [(192, 8, 100000), (192, 4, 200000), (256, 8, 75000), (256, 4, 150000), (128, 8, 150000), (192, 8, 400000)]:
    out.append(capi.debug_i8_peak(n_cols, pc, iters, reps=3))
print(json.dumps(out))
PY
cat gpurun_out/r02a_i8_peak.json; tail -3 gpurun_out/r02a_i8_peak.err
