#!/bin/bash
# round 2, call V (8 GPUs): the final (fp4) kernels -- bench at N=8 with the strong record, the UKB-scale genome again (8 oracle slices)
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/r02v_bench_n8.json 2> gpurun_out/r02v_bench_n8.err; echo "bench n8 rc $?"
python -c "
import json; d=json.load(open('gpurun_out/r02v_bench_n8.json')); print(d['value'], d['ms_per_step'], d['e2e']['ms_per_step'], d['roofline']['e2e']['h2d_gbs_n_way'], d['roofline']['frac'], d['roofline']['frac_of_int8_peak']); print(json.dumps(d.get('strong'))[:1400])"
tail -2 gpurun_out/r02v_bench_n8.err
NLDSC_UKB_ORACLE_CHROMS=1,5,9,13,17,20,21,22 timeout 900 python scripts/ukb_genome.py 8 1200000 > gpurun_out/r02v_ukb.log 2> gpurun_out/r02v_ukb.err; echo "ukb rc $?"; tail -c 2300 gpurun_out/r02v_ukb.log; tail -2 gpurun_out/r02v_ukb.err
cp gpurun_out/r02_ukb_genome_8gpu.json gpurun_out/r02v_ukb_genome_8gpu.json
