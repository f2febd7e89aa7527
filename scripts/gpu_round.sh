# One gpurun call: GPU tests, smoke, bench (both arms), then the two ncu passes on the bench command.
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
nproc >> gpurun_out/gpu.txt; lscpu | grep -E "Model name|Socket|Core|Thread" >> gpurun_out/gpu.txt
: > gpurun_out/rc.txt
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/t_all.log 2>&1; echo "all rc=$?" >> gpurun_out/rc.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/rc.txt
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/rc.txt
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.log 2> gpurun_out/bench_ref.err; echo "bench_ref rc=$?" >> gpurun_out/rc.txt
if [ "${NCU:-1}" = "1" ]; then
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/plain.log 2>&1 && \
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -k "regex:band_|unpack_|count_kernel|stats_kernel|finalize_|zero_rows|windows_kernel|copy16_kernel|peak_probe" -c 600 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu_launches rc=$?" >> gpurun_out/rc.txt
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:band_tc -s 39 -c 13 -f -o gpurun_out/band_tc $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu_full rc=$?" >> gpurun_out/rc.txt
fi
cat gpurun_out/rc.txt; tail -5 gpurun_out/t_all.log; cat gpurun_out/smoke.log | tail -2; cat gpurun_out/bench.log; tail -5 gpurun_out/bench.err; cat gpurun_out/bench_ref.log
