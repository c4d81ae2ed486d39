#!/bin/bash
# GPU call 8: how ms_kernel takes fewer resident blocks per SM (dynamic shared memory padding): 10 / 9 / 8 / 7 / 6 blocks
mkdir -p gpurun_out
OUT=gpurun_out/r2_ab_ms_occupancy.txt
: > $OUT
for pad in 0 12000 15000 19000 25000; do
  NTC_MS_PAD_SMEM=$pad timeout 300 python tools/ab_ms.py pad$pad 2>&1 | tail -2 >> $OUT
done
cat $OUT
