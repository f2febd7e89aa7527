# 2-GPU check: GPU tests (incl. the NLDSC_GPUS=2 sharded path) and the torchrun bench at N=2, both arms.
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/t2_all.log 2>&1; echo "tests rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_n2.log 2> gpurun_out/bench_n2.err; echo "bench2 rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/bench_ref_n2.log 2> gpurun_out/bench_ref_n2.err; echo "bench_ref2 rc=$?"
tail -4 gpurun_out/t2_all.log; cat gpurun_out/bench_n2.log; tail -3 gpurun_out/bench_n2.err; cat gpurun_out/bench_ref_n2.log
