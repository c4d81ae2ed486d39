#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/gpu_tests.log 2>&1; echo "gpu tests rc=$?"; tail -3 gpurun_out/gpu_tests.log
python tools/pageable_probe.py > gpurun_out/pageable_probe2.json 2> gpurun_out/pageable_probe2.err; echo rc=$?; cat gpurun_out/pageable_probe2.json; tail -2 gpurun_out/pageable_probe2.err
NTC_HOST_THREADS=0 python tools/pageable_probe.py 2000000 > /dev/null 2>&1; echo "threads0 rc=$?"
