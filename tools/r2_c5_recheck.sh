#!/bin/bash
mkdir -p gpurun_out
OUT=gpurun_out/c5_recheck.txt; : > $OUT
timeout 300 python tools/ab_ms.py c5_auto_a 250000 10000 50000 2>&1 | tail -1 >> $OUT
timeout 300 python tools/ab_ms.py c3_auto 2>&1 | tail -1 >> $OUT
timeout 300 python tools/ab_ms.py c4_auto 12500000 150 10000 12000000 2>&1 | tail -1 >> $OUT
timeout 300 python tools/ab_ms.py c5_auto_b 250000 10000 50000 2>&1 | tail -1 >> $OUT
timeout 300 python tools/ab_ms.py l2k_auto 1000000 2000 10000 2>&1 | tail -1 >> $OUT
cat $OUT
