#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/ab_ms.py c4_small 4000000 150 10000 12000000 2>&1 | tail -1 > gpurun_out/c4_small.txt || exit 1
ncu --set full --clock-control none --import-source on -k regex:'ms_kernel' -s 3 -c 1 -f -o gpurun_out/prof_c4_ms python tools/ab_ms.py c4_small 4000000 150 10000 12000000 > gpurun_out/ncu_c4.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/c4_small.txt
