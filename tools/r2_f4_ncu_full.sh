#!/bin/bash
mkdir -p gpurun_out
python tools/index_build_once.py 5000000 91 > gpurun_out/f4_once.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:'radix_hist_kernel|radix_scatter_kernel' -s 8 -c 2 -f -o gpurun_out/prof_f4 python tools/index_build_once.py 5000000 91 > gpurun_out/ncu_f4.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/prof_f4.ncu-rep
