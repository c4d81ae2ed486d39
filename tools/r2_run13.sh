#!/bin/bash
# GPU call 13: 16-bit LCS side array against the 32-bit one: C3 (5 Mbp) and the C4 shape (12 Mbp genome, one 8-GPU shard)
mkdir -p gpurun_out
OUT=gpurun_out/r2_ab_lcs16.txt
: > $OUT
run() { local label=$1 lib=$2; shift 2; env "$@" NTC_LIB_PATH=$lib timeout 400 python tools/ab_ms.py $label $ABARGS 2>&1 | tail -1 >> $OUT; }
D=$PWD/ntcomp_b200/libntcomp_b200.so
V=$PWD/build/lcs16.so
ABARGS="" run c3_lcs32 $D X=1
ABARGS="" run c3_lcs16 $V X=1
ABARGS="12500000 150 10000 12000000" run c4_lcs32 $D X=1
ABARGS="12500000 150 10000 12000000" run c4_lcs16 $V X=1
ABARGS="250000 10000 50000" run c5_lcs32 $D X=1
ABARGS="250000 10000 50000" run c5_lcs16 $V X=1
cat $OUT
