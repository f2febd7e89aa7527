#!/bin/bash
# round 2, call P (4 GPUs): N-way H2D ceiling with default / write-combined pinned memory; genome driver (native .bim) at 1 / 2 / 4 GPUs
# on configs[2] and on a 22-chromosome UKB-shaped cut read from .bed files
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
nproc > gpurun_out/r02p_box.txt; free -g >> gpurun_out/r02p_box.txt
timeout 300 python - > gpurun_out/r02p_h2d_probe.json 2> gpurun_out/r02p_h2d_probe.err <<'PY'
import json
from nldsc_b200 import capi
out = {}
for n in (1, 2, 4):
    for wc in (False, True):
        out[f"{n}gpu_{'write_combined' if wc else 'default'}"] = capi.debug_h2d_probe(1 << 30, n, wc, reps=4)
print(json.dumps(out))
PY
cat gpurun_out/r02p_h2d_probe.json; tail -2 gpurun_out/r02p_h2d_probe.err
timeout 600 python scripts/cfg3_genome.py 1,2,4 100000 > gpurun_out/r02p_cfg3.log 2> gpurun_out/r02p_cfg3.err; echo "cfg3 rc $?"
python - <<'PY'
import json, shutil
for g in (1, 2, 4):
    try:
        d = json.load(open(f"gpurun_out/cfg3_genome_{g}gpu.json"))
        shutil.copy(f"gpurun_out/cfg3_genome_{g}gpu.json", f"gpurun_out/r02p_cfg3_genome_{g}gpu.json")
        print(g, "gpus: wall", d["wall_s"], "phases", d["phases_last_run"], "busy", d["per_gpu_busy_s_last_run"])
    except Exception as e:
        print(g, e)
PY
NLDSC_GENOME=ukb timeout 900 python scripts/cfg3_genome.py 1,4 5000 > gpurun_out/r02p_cfg4cut.log 2> gpurun_out/r02p_cfg4cut.err; echo "cfg4cut rc $?"
python - <<'PY'
import json, shutil
for g in (1, 4):
    try:
        d = json.load(open(f"gpurun_out/cfg4_genome_{g}gpu.json"))
        shutil.copy(f"gpurun_out/cfg4_genome_{g}gpu.json", f"gpurun_out/r02p_cfg4cut_genome_{g}gpu.json")
        print(g, "gpus:", d["config"], "wall", d["wall_s"], "phases", d["phases_last_run"], "busy", d["per_gpu_busy_s_last_run"])
    except Exception as e:
        print(g, e)
PY
tail -3 gpurun_out/r02p_cfg4cut.err
