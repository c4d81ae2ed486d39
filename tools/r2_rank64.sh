#!/bin/bash
mkdir -p gpurun_out
OUT=gpurun_out/ab_rank64.txt; : > $OUT
run() { local label=$1 lib=$2; shift 2; NTC_LIB_PATH=$lib timeout 300 python tools/ab_ms.py $label "$@" 2>&1 | tail -1 >> $OUT; }
D=$PWD/ntcomp_b200/libntcomp_b200.so; V=$PWD/build/rank64.so
run c3_rank32 $D;                      run c3_rank64 $V
run c4_rank32 $D 12500000 150 10000 12000000; run c4_rank64 $V 12500000 150 10000 12000000
run c5_rank32 $D 250000 10000 50000;   run c5_rank64 $V 250000 10000 50000
run c2k31_rank32 $D 1000000 150 10000; run c2k31_rank64 $V 1000000 150 10000
cat $OUT
