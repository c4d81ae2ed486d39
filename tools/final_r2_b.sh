#!/bin/bash
# round-2 evidence, call B: ncu --set full of the dominant kernels in the bench command (after the same command ran clean)
TAG=${1:-r2}
mkdir -p gpurun_out
python bench.py --config C5 --cpu-budget 0 --steps 3 > gpurun_out/bench_c5_$TAG.json 2> gpurun_out/bench_c5_$TAG.err; echo "C5 rc=$?"
python bench.py --steps 1 --warmup 3 --cpu-budget 0 --no-e2e > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'ms_kernel|walk_kernel|emit_kernel|pack2bit_kernel' -s 16 -c 4 -f -o gpurun_out/prof_$TAG python bench.py --steps 1 --warmup 3 --cpu-budget 0 --no-e2e > gpurun_out/ncu_full_$TAG.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/prof_$TAG.ncu-rep; tail -3 gpurun_out/ncu_full_$TAG.log
