#!/bin/bash
# round 2, call Z (1 GPU): the allocation-free entry point alone in a fresh process, then the whole GPU suite (which ends with it),
# then the bench and its reference arm on the final binary
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r02z_no_alloc_alone.txt 2>&1; echo "no_alloc alone rc $?"; tail -5 gpurun_out/r02z_no_alloc_alone.txt | cut -c1-300
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r02z_gpu_tests.txt 2>&1; echo "suite rc $?"; tail -5 gpurun_out/r02z_gpu_tests.txt | cut -c1-300
timeout 400 python bench.py --steps 20 --warmup 3 > gpurun_out/r02z_bench.json 2> gpurun_out/r02z_bench.err; echo "bench rc $?"; cut -c1-1500 gpurun_out/r02z_bench.json
timeout 400 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02z_bench_reference.json 2> gpurun_out/r02z_bench_reference.err; echo "ref rc $?"; cut -c1-600 gpurun_out/r02z_bench_reference.json
