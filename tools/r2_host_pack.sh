#!/bin/bash
mkdir -p gpurun_out
cd tools/host_pack_probe && g++ -O2 -std=c++17 -pthread -o /tmp/probe probe.cpp ../../ntcomp_b200/csrc/host_pack.cpp && cd ../..
OUT=gpurun_out/host_pack_probe.txt; : > $OUT
nproc >> $OUT; grep -m1 "model name" /proc/cpuinfo >> $OUT
for t in 1 2 4 8 16; do /tmp/probe $t 1500000000 | tail -2 | head -1 >> $OUT; done
cat $OUT
