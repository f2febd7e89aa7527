#!/bin/bash
# round 2, call L (2 GPUs): final code -- full GPU suite on a multi-GPU box, genome driver with the native .bim reader at 1 / 2 GPUs,
# bench --gpus 2
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r02l_gpu_tests_2gpu.txt 2>&1; echo "pytest rc $?" >> gpurun_out/r02l_gpu_tests_2gpu.txt
tail -4 gpurun_out/r02l_gpu_tests_2gpu.txt
timeout 600 python scripts/cfg3_genome.py 1,2 100000 > gpurun_out/r02l_cfg3.log 2> gpurun_out/r02l_cfg3.err; echo "cfg3 rc $?"
python - <<'PY'
import json, shutil
for g in (1, 2):
    try:
        d = json.load(open(f"gpurun_out/cfg3_genome_{g}gpu.json"))
        shutil.copy(f"gpurun_out/cfg3_genome_{g}gpu.json", f"gpurun_out/r02l_cfg3_genome_{g}gpu.json")
        print(g, "gpus: wall", d["wall_s"], "phases", d["phases_last_run"], "busy", d["per_gpu_busy_s_last_run"])
    except Exception as e:
        print(g, e)
PY
tail -3 gpurun_out/r02l_cfg3.err
NLDSC_BENCH_STRONG_SNPS=120000 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29520 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/r02l_bench_n2.json 2> gpurun_out/r02l_bench_n2.err; echo "bench n2 rc $?"
python -c "
import json; d=json.load(open('gpurun_out/r02l_bench_n2.json')); print(d['value'], d['ms_per_step'], d['e2e']['ms_per_step']); print(json.dumps(d.get('strong'))[:900])"
