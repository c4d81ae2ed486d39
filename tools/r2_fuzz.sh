#!/bin/bash
mkdir -p gpurun_out
for seed in ${SEEDS:-1 2 3}; do timeout 400 python tools/fuzz_parity.py ${SECS:-75} $seed 2> gpurun_out/fuzz_$seed.err | tail -1 | tee -a gpurun_out/fuzz_parity.jsonl; echo "seed $seed rc=${PIPESTATUS[0]}"; done
tail -3 gpurun_out/fuzz_1.err
