#!/bin/bash
# GPU call 10: resident blocks per SM with the L1 kept large (stage buffer 608 words)
mkdir -p gpurun_out
OUT=gpurun_out/r2_ab_ms_blocks.txt
: > $OUT
run() { local label=$1 lib=$2; shift 2; env "$@" NTC_LIB_PATH=$lib timeout 300 python tools/ab_ms.py $label 2>&1 | tail -1 >> $OUT; }
run default $PWD/ntcomp_b200/libntcomp_b200.so X=1
for v in b8 b9 b11 b12; do run $v $PWD/build/$v.so X=1; run ${v}_carve44 $PWD/build/$v.so NTC_MS_CARVEOUT=44; done
cat $OUT
