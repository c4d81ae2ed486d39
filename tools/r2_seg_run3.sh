#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_parity.py -x -q -m gpu > gpurun_out/seg_tests3.log 2>&1; echo "parity rc=$?"; tail -2 gpurun_out/seg_tests3.log
OUT=gpurun_out/ab_seg_parallel3.txt; : > $OUT
for shape in "1 5000000 10000" "16 1000000 10000" "2000 100000 50000" "20000 10000 50000"; do
  set -- $shape
  NTC_SEG_SERIAL=0 timeout 300 python tools/ab_ms.py segments_adaptiveS_$1x$2 $1 $2 $3 2>&1 | tail -1 >> $OUT
done
NTC_SEG_SERIAL=0 NTC_SEG_LEN=4096 timeout 300 python tools/ab_ms.py segments_S4096_1x5000000 1 5000000 10000 2>&1 | tail -1 >> $OUT
cat $OUT
