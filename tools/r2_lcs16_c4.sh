#!/bin/bash
mkdir -p gpurun_out
OUT=gpurun_out/ab_lcs16_rank64.txt; : > $OUT
V=$PWD/build/lcs16.so
timeout 300 python tools/ab_ms.py c4_rank64_lcs32 12500000 150 10000 12000000 2>&1 | tail -1 >> $OUT
NTC_LIB_PATH=$V timeout 300 python tools/ab_ms.py c4_rank64_lcs16 12500000 150 10000 12000000 2>&1 | tail -1 >> $OUT
timeout 300 python tools/ab_ms.py g24_rank64_lcs32 12500000 150 10000 24000000 2>&1 | tail -1 >> $OUT
NTC_LIB_PATH=$V timeout 300 python tools/ab_ms.py g24_rank64_lcs16 12500000 150 10000 24000000 2>&1 | tail -1 >> $OUT
cat $OUT
