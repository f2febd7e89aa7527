set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
: > gpurun_out/rc.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/t_all.log 2>&1; echo "all rc=$?" >> gpurun_out/rc.txt
timeout 300 python scripts/e2e_trace.py > gpurun_out/trace.log 2>&1; echo "trace rc=$?" >> gpurun_out/rc.txt
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt; tail -15 gpurun_out/t_all.log; tail -14 gpurun_out/trace.log; cat gpurun_out/bench.log; tail -5 gpurun_out/bench.err
