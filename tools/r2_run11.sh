#!/bin/bash
# GPU call 11: dynamic stage buffer + 12-block instantiation: GPU tests, then A/B against the forced 10-block kernel
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
OUT=gpurun_out/r2_ab_ms_blocks_adopted.txt
: > $OUT
run() { local label=$1; shift; env "$@" timeout 300 python tools/ab_ms.py $label $ABARGS 2>&1 | tail -1 >> $OUT; }
ABARGS="" run c3_default X=1
ABARGS="" run c3_force10 NTC_MS_BLOCKS=10 NTC_MS_STAGE_WORDS=1024
ABARGS="6000000 250 10000" run r250_default X=1
ABARGS="6000000 250 10000" run r250_force12 NTC_MS_BLOCKS=12
ABARGS="15000000 100 10000" run r100_default X=1
ABARGS="15000000 100 10000" run r100_force10 NTC_MS_BLOCKS=10 NTC_MS_STAGE_WORDS=1024
ABARGS="250000 10000 50000" run c5_default X=1
cat $OUT
