#!/bin/bash
# round 2, GPU call 1: GPU suite with the full-size configs, bench line, walk_kernel store-policy A/B (C3 and C4 shapes)
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q --durations=12 2>&1 | tail -25 > gpurun_out/r2_gpu_tests_1.log; tail -4 gpurun_out/r2_gpu_tests_1.log
python bench.py > gpurun_out/r2_bench_c3_a.json 2> gpurun_out/r2_bench_c3_a.err; echo "bench rc=$?"; tail -3 gpurun_out/r2_bench_c3_a.err
ABARGS="10000000 150 10000 5000000" tools/ab_run.sh walk_nostream
ABARGS="12517376 150 10000 12000000" tools/ab_run.sh walk_nostream
head -c 600 gpurun_out/r2_bench_c3_a.json
