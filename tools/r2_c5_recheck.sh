#!/bin/bash
mkdir -p gpurun_out
OUT=gpurun_out/c5_recheck.txt; : > $OUT
timeout 300 python tools/ab_ms.py c5_auto_a 250000 10000 50000 2>&1 | tail -1 >> $OUT
NTC_RANK_NODES=32 timeout 300 python tools/ab_ms.py c5_force32 250000 10000 50000 2>&1 | tail -1 >> $OUT
timeout 300 python tools/ab_ms.py c5_auto_b 250000 10000 50000 2>&1 | tail -1 >> $OUT
cat $OUT
