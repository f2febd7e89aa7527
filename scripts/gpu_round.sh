set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
nproc >> gpurun_out/gpu.txt; lscpu | grep -E "Model name|Socket|Core|Thread" >> gpurun_out/gpu.txt
timeout 600 python -m pytest tests -m gpu -x -q -k "unpack or simt" > gpurun_out/t_simt.log 2>&1; echo "simt rc=$?" >> gpurun_out/rc.txt
timeout 600 python -m pytest tests -m gpu -x -q -k "cross_counts and tc" > gpurun_out/t_tc_cc.log 2>&1; echo "tc_cc rc=$?" >> gpurun_out/rc.txt
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/t_all.log 2>&1; echo "all rc=$?" >> gpurun_out/rc.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/rc.txt
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt; tail -5 gpurun_out/t_all.log; cat gpurun_out/bench.log; tail -5 gpurun_out/bench.err
