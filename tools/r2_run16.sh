#!/bin/bash
# GPU call 16: walk_kernel with a persisting L2 window over the jump table (set-aside x hit ratio), C3 and the C4 shape
mkdir -p gpurun_out
OUT=gpurun_out/r2_ab_walk_window.txt
: > $OUT
run() { local label=$1; shift; env "$@" timeout 400 python tools/ab_ms.py $label $ABARGS 2>&1 | tail -1 >> $OUT; }
ABARGS="" run c3_default X=1
ABARGS="" run c3_set32_hit40 NTC_L2_PERSIST_MB=32 NTC_WALK_WINDOW=40
ABARGS="" run c3_set64_hit80 NTC_L2_PERSIST_MB=64 NTC_WALK_WINDOW=80
ABARGS="" run c3_set79_hit100 NTC_L2_PERSIST_MB=79 NTC_WALK_WINDOW=100
ABARGS="12500000 150 10000 12000000" run c4_default X=1
ABARGS="12500000 150 10000 12000000" run c4_set64_hit33 NTC_L2_PERSIST_MB=64 NTC_WALK_WINDOW=33
cat $OUT
