#!/bin/bash
mkdir -p gpurun_out
OUT=gpurun_out/ab_rank64b.txt; : > $OUT
V=$PWD/build/rank64.so
NTC_MS_BLOCKS=10 NTC_LIB_PATH=$V timeout 300 python tools/ab_ms.py c4_rank64_blocks10 12500000 150 10000 12000000 2>&1 | tail -1 >> $OUT
NTC_MS_BLOCKS=10 NTC_LIB_PATH=$V timeout 300 python tools/ab_ms.py c3_rank64_blocks10 2>&1 | tail -1 >> $OUT
NTC_LIB_PATH=$V timeout 300 python tools/ab_ms.py g24_rank64 12500000 150 10000 24000000 2>&1 | tail -1 >> $OUT
timeout 300 python tools/ab_ms.py g24_rank32 12500000 150 10000 24000000 2>&1 | tail -1 >> $OUT
cat $OUT
