#!/bin/bash
# segment-parallel segmentation of long reads: GPU parity tests + A/B on C5 (NTC_SEG_SERIAL=1 = the one-thread-per-read count_kernel)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_parity.py -x -q -m gpu > gpurun_out/seg_tests.log 2>&1; echo "parity rc=$?" >> gpurun_out/seg_tests.log
timeout 900 python -m pytest tests/test_full_size.py -x -q -m gpu -k "c5 or C5" >> gpurun_out/seg_tests.log 2>&1; echo "full-size c5 rc=$?" >> gpurun_out/seg_tests.log
tail -4 gpurun_out/seg_tests.log
OUT=gpurun_out/ab_seg_parallel.txt; : > $OUT
NTC_SEG_SERIAL=1 timeout 300 python tools/ab_ms.py c5_serial 250000 10000 50000 2>&1 | tail -1 >> $OUT
timeout 300 python tools/ab_ms.py c5_segments 250000 10000 50000 2>&1 | tail -1 >> $OUT
NTC_SEG_SERIAL=1 timeout 300 python tools/ab_ms.py l2k_serial 1000000 2000 10000 2>&1 | tail -1 >> $OUT
timeout 300 python tools/ab_ms.py l2k_segments 1000000 2000 10000 2>&1 | tail -1 >> $OUT
timeout 300 python tools/ab_ms.py c3_default 2>&1 | tail -1 >> $OUT
cat $OUT
