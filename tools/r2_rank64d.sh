#!/bin/bash
mkdir -p gpurun_out
OUT=gpurun_out/ab_rank64d.txt; : > $OUT
timeout 300 python tools/ab_ms.py c4_rank64_halves 12500000 150 10000 12000000 2>&1 | tail -1 >> $OUT
timeout 300 python tools/ab_ms.py g24_rank64_halves 12500000 150 10000 24000000 2>&1 | tail -1 >> $OUT
NTC_RANK_NODES=64 timeout 300 python tools/ab_ms.py c3_rank64_halves 2>&1 | tail -1 >> $OUT
NTC_RANK_NODES=64 timeout 300 python tools/ab_ms.py g8_rank64_halves 12500000 150 10000 8000000 2>&1 | tail -1 >> $OUT
NTC_RANK_NODES=64 timeout 300 python tools/ab_ms.py g10_rank64_halves 12500000 150 10000 10000000 2>&1 | tail -1 >> $OUT
timeout 300 python tools/ab_ms.py c3_default 2>&1 | tail -1 >> $OUT
cat $OUT
