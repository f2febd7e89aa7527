# Round-end dress rehearsal: what the driver runs, in its order.
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
: > gpurun_out/rc.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/t_final.log 2>&1; echo "tests rc=$?" >> gpurun_out/rc.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke_final.log 2>&1; echo "smoke rc=$?" >> gpurun_out/rc.txt
timeout 600 python bench.py --impl reference > gpurun_out/ref_final.json 2> gpurun_out/ref_final.err; echo "ref rc=$?" >> gpurun_out/rc.txt
timeout 600 python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err; echo "bench rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt; tail -4 gpurun_out/t_final.log; tail -2 gpurun_out/smoke_final.log; cat gpurun_out/ref_final.json gpurun_out/bench_final.json
