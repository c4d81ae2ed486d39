#!/bin/bash
# round-2 evidence, call C: the fused column codec (row f1): plain run, then ncu --set full of its kernels
TAG=${1:-r2}
mkdir -p gpurun_out
python tools/columns_probe.py 16 > gpurun_out/columns_probe_$TAG.txt 2>&1 && cat gpurun_out/columns_probe_$TAG.txt && \
ncu --set full --clock-control none --import-source on -k regex:'split_kernel|blk_stats_kernel|lit_pack_kernel|seg_finish_kernel|pack_words_kernel|cw_table_kernel|cw_resolve_kernel|cw_extract_kernel|zip_kernel|blk_trim_kernel' -s 20 -c 10 -f -o gpurun_out/prof_columns_$TAG python tools/columns_probe.py 16 > gpurun_out/ncu_columns_$TAG.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/prof_columns_$TAG.ncu-rep; tail -3 gpurun_out/ncu_columns_$TAG.log
