#!/bin/bash
# one gpurun call: A/B of kernel-variant builds under build/ (development aid)
#   ABARGS="n_reads read_len err_ppm G" tools/ab_run.sh variant ...     (default = the in-tree library)
mkdir -p gpurun_out
OUT=gpurun_out/ab_$(date +%H%M%S).txt
run() { # label lib env...
  local label=$1 lib=$2; shift 2
  env "$@" NTC_LIB_PATH=$lib timeout 300 python tools/ab_ms.py $label $ABARGS 2>&1 | tail -2 >> $OUT
}
run default $PWD/ntcomp_b200/libntcomp_b200.so X=1
for v in "$@"; do run $v $PWD/build/$v.so X=1; done
cat $OUT
