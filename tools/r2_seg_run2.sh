#!/bin/bash
mkdir -p gpurun_out
OUT=gpurun_out/ab_seg_parallel2.txt; : > $OUT
for shape in "1 5000000 10000" "16 1000000 10000" "2000 100000 50000" "20000 10000 50000"; do
  set -- $shape
  NTC_SEG_SERIAL=1 timeout 300 python tools/ab_ms.py serial_$1x$2 $1 $2 $3 2>&1 | tail -1 >> $OUT
  NTC_SEG_SERIAL=0 timeout 300 python tools/ab_ms.py segments_$1x$2 $1 $2 $3 2>&1 | tail -1 >> $OUT
done
timeout 300 python tools/ab_ms.py c3_default 2>&1 | tail -1 >> $OUT
timeout 300 python tools/ab_ms.py c5_default 250000 10000 50000 2>&1 | tail -1 >> $OUT
cat $OUT
