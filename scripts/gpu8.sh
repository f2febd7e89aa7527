# 8-GPU preview of the driver's scaling run: torchrun bench at N=8 (both arms) and N=4.
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
for n in 8 4; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/bench_n$n.log 2> gpurun_out/bench_n$n.err; echo "bench$n rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/bench_n$n.log')); print('N=$n value %.4g ms/step %.3f e2e %.4g (%.3f ms) frac %.3f clocks %s' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['roofline']['frac'], d['clocks']))"
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29519 bench.py --impl reference --gpus 8 --steps 2 --warmup 1 > gpurun_out/bench_ref_n8.log 2> gpurun_out/bench_ref_n8.err; echo "bench_ref8 rc=$?"; cut -c1-200 gpurun_out/bench_ref_n8.log; tail -3 gpurun_out/bench_n8.err
