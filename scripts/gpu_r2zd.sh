#!/bin/bash
# round 2, call ZD (2 GPUs): the driver's launch of the bench at N = 2 on the final binary (weak record + strong record with bitwise check)
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/r02zd_bench_n2.json 2> gpurun_out/r02zd_bench_n2.err; echo "bench n2 rc $?"
python -c "
import json; d=json.load(open('gpurun_out/r02zd_bench_n2.json')); r=d['roofline']
print('value', d['value'], 'step', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], 'strong', d.get('strong'))" | cut -c1-900
