#!/bin/bash
mkdir -p gpurun_out
OUT=gpurun_out/ab_c5_blocks.txt; : > $OUT
timeout 300 python tools/ab_ms.py c5_default 250000 10000 50000 2>&1 | tail -1 >> $OUT
NTC_MS_STAGE_WORDS=0 NTC_MS_BLOCKS=12 timeout 300 python tools/ab_ms.py c5_stage0_blocks12 250000 10000 50000 2>&1 | tail -1 >> $OUT
NTC_MS_STAGE_WORDS=0 NTC_MS_BLOCKS=10 timeout 300 python tools/ab_ms.py c5_stage0_blocks10 250000 10000 50000 2>&1 | tail -1 >> $OUT
NTC_MS_BLOCKS=12 timeout 300 python tools/ab_ms.py c5_stage1024_blocks12 250000 10000 50000 2>&1 | tail -1 >> $OUT
NTC_TILE_LEN=1024 timeout 300 python tools/ab_ms.py c5_T1024 250000 10000 50000 2>&1 | tail -1 >> $OUT
NTC_TILE_LEN=3456 timeout 300 python tools/ab_ms.py c5_T3456 250000 10000 50000 2>&1 | tail -1 >> $OUT
NTC_MS_STAGE_WORDS=0 NTC_MS_BLOCKS=12 NTC_TILE_LEN=1024 timeout 300 python tools/ab_ms.py c5_stage0_blocks12_T1024 250000 10000 50000 2>&1 | tail -1 >> $OUT
cat $OUT
