#!/bin/bash
mkdir -p gpurun_out
OUT=gpurun_out/ab_ms_threads.txt; : > $OUT
run() { local label=$1 lib=$2; shift 2; NTC_LIB_PATH=$lib timeout 300 python tools/ab_ms.py $label "$@" 2>&1 | tail -1 >> $OUT; }
D=$PWD/ntcomp_b200/libntcomp_b200.so; V=$PWD/build/ms64.so
run c3_t128 $D;                      run c3_t64 $V
run c4_t128 $D 12500000 150 10000 12000000; run c4_t64 $V 12500000 150 10000 12000000
run c5_t128 $D 250000 10000 50000;   run c5_t64 $V 250000 10000 50000
run c2_t128 $D 1000000 150 10000; run c2_t64 $V 1000000 150 10000
cat $OUT
