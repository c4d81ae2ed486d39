#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_parity.py tests/test_full_size.py -x -q -m gpu > gpurun_out/rank64_tests.log 2>&1; echo "tests rc=$?"; tail -2 gpurun_out/rank64_tests.log
OUT=gpurun_out/ab_rank64c.txt; : > $OUT
timeout 300 python tools/ab_ms.py c3_auto 2>&1 | tail -1 >> $OUT
timeout 300 python tools/ab_ms.py c4_auto 12500000 150 10000 12000000 2>&1 | tail -1 >> $OUT
timeout 300 python tools/ab_ms.py c5_auto 250000 10000 50000 2>&1 | tail -1 >> $OUT
for G in 8000000 10000000; do
NTC_RANK_NODES=32 timeout 300 python tools/ab_ms.py g${G}_rank32 12500000 150 10000 $G 2>&1 | tail -1 >> $OUT
NTC_RANK_NODES=64 timeout 300 python tools/ab_ms.py g${G}_rank64 12500000 150 10000 $G 2>&1 | tail -1 >> $OUT
done
cat $OUT
