#!/bin/bash
# round-end evidence run on one B200: GPU tests, bench (both arms), ncu launch list + full capture of the dominant kernels
TAG=${1:-v12}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee gpurun_out/gpu_tests_$TAG.log
python bench.py > gpurun_out/bench_c3_$TAG.json 2> gpurun_out/bench_c3_$TAG.err; echo "bench rc=$?"
python bench.py --impl reference > gpurun_out/bench_c3_reference_arm_$TAG.json 2> gpurun_out/bench_ref_$TAG.err; echo "ref rc=$?"
python bench.py --steps 2 --warmup 3 --cpu-budget 0 --no-e2e > gpurun_out/bench_plain_$TAG.json 2>/dev/null && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_$TAG.csv python bench.py --steps 2 --warmup 3 --cpu-budget 0 --no-e2e > gpurun_out/ncu_launch_$TAG.log 2>&1
python bench.py --steps 1 --warmup 3 --cpu-budget 0 --no-e2e > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'ms_kernel|walk_kernel|emit_kernel|pack2bit_kernel' -s 16 -c 4 -f -o gpurun_out/prof_$TAG python bench.py --steps 1 --warmup 3 --cpu-budget 0 --no-e2e > gpurun_out/ncu_full_$TAG.log 2>&1
cat gpurun_out/bench_c3_$TAG.json | head -c 1500; echo; tail -3 gpurun_out/bench_c3_$TAG.err
