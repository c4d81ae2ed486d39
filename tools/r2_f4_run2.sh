#!/bin/bash
mkdir -p gpurun_out
python tools/index_build_once.py 5000000 91 > gpurun_out/f4_once.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/f4_launches.csv python tools/index_build_once.py 5000000 91 > gpurun_out/f4_ncu.log 2>&1
echo "ncu rc=$?"
tail -2 gpurun_out/f4_once.log
