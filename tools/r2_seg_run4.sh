#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_parity.py tests/test_full_size.py -x -q -m gpu 2>&1 | tail -2
OUT=gpurun_out/ab_seg_parallel4.txt; : > $OUT
for shape in "250000 10000 50000" "1000000 2000 10000" "20000 10000 50000" "2000 100000 50000" "1 5000000 10000"; do
  set -- $shape
  NTC_SEG_SERIAL=1 timeout 300 python tools/ab_ms.py serial_$1x$2 $1 $2 $3 2>&1 | tail -1 >> $OUT
  NTC_SEG_SERIAL=0 timeout 300 python tools/ab_ms.py segslots_$1x$2 $1 $2 $3 2>&1 | tail -1 >> $OUT
done
cat $OUT
