#!/bin/bash
# round 2, call H (8 GPUs): the north_star Target (UKB-scale genome end to end), bench at N=8 and N=4 (strong records),
# the configs[2] genome through the .bed driver at 1/2/4/8 GPUs
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name,clocks.max.sm --format=csv > gpurun_out/r02h_box.txt; nproc >> gpurun_out/r02h_box.txt; free -g >> gpurun_out/r02h_box.txt; df -h /tmp >> gpurun_out/r02h_box.txt; lscpu | grep -E "Model name|Socket|NUMA" >> gpurun_out/r02h_box.txt
nvidia-smi topo -m >> gpurun_out/r02h_box.txt 2>&1
timeout 1200 python scripts/ukb_genome.py 8 1200000 > gpurun_out/r02h_ukb.log 2> gpurun_out/r02h_ukb.err; echo "ukb rc $?"; tail -c 2500 gpurun_out/r02h_ukb.log; tail -3 gpurun_out/r02h_ukb.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/r02h_bench_n8.json 2> gpurun_out/r02h_bench_n8.err; echo "bench n8 rc $?"
python -c "
import json; d=json.load(open('gpurun_out/r02h_bench_n8.json')); print(d['value'], d['ms_per_step'], d['e2e']['ms_per_step'], d['roofline']['e2e']); print(json.dumps(d.get('strong')))"
tail -3 gpurun_out/r02h_bench_n8.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus 4 --steps 20 --warmup 3 > gpurun_out/r02h_bench_n4.json 2> gpurun_out/r02h_bench_n4.err; echo "bench n4 rc $?"
python -c "
import json; d=json.load(open('gpurun_out/r02h_bench_n4.json')); print(d['value'], d['ms_per_step'], d['e2e']['ms_per_step'], d['roofline']['e2e']); print(json.dumps(d.get('strong')))"
timeout 900 python scripts/cfg3_genome.py 1,2,4,8 100000 > gpurun_out/r02h_cfg3.log 2> gpurun_out/r02h_cfg3.err; echo "cfg3 rc $?"
python - <<'PY'
import json
for g in (1, 2, 4, 8):
    try:
        d = json.load(open(f"gpurun_out/cfg3_genome_{g}gpu.json"))
        print(g, "gpus: wall", d["wall_s"], "phases", d["phases_last_run"], "busy", d["per_gpu_busy_s_last_run"])
    except Exception as e:
        print(g, e)
PY
tail -3 gpurun_out/r02h_cfg3.err
