#!/bin/bash
# GPU call 9: ms_kernel against the L1 / shared-memory split (the smem-padding probe pointed at L1 capacity, not occupancy)
mkdir -p gpurun_out
OUT=gpurun_out/r2_ab_ms_l1.txt
: > $OUT
run() { local label=$1 lib=$2; shift 2; env "$@" NTC_LIB_PATH=$lib timeout 300 python tools/ab_ms.py $label 2>&1 | tail -1 >> $OUT; }
D=$PWD/ntcomp_b200/libntcomp_b200.so
run default $D X=1
run default_carve58 $D NTC_MS_CARVEOUT=58
run default_carve100 $D NTC_MS_CARVEOUT=100
run stage608 $PWD/build/stage608.so X=1
run stage608_carve44 $PWD/build/stage608.so NTC_MS_CARVEOUT=44
run stage608_carve58 $PWD/build/stage608.so NTC_MS_CARVEOUT=58
run stage16 $PWD/build/stage16.so X=1
run stage16_carve28 $PWD/build/stage16.so NTC_MS_CARVEOUT=28
run stage16_carve44 $PWD/build/stage16.so NTC_MS_CARVEOUT=44
cat $OUT
