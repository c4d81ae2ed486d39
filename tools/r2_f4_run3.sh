#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_index_build.py -x -q -m gpu > gpurun_out/f4_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/f4_tests.log; tail -3 gpurun_out/f4_tests.log
timeout 600 python tools/index_build_bench.py gpurun_out/index_build_bench.jsonl > gpurun_out/f4_bench.log 2>&1
echo "bench rc=$?"; cat gpurun_out/index_build_bench.jsonl; tail -3 gpurun_out/f4_bench.log
python tools/index_build_once.py 5000000 91 > gpurun_out/f4_once.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/f4_launches.csv python tools/index_build_once.py 5000000 91 > gpurun_out/f4_ncu.log 2>&1
echo "ncu rc=$?"
