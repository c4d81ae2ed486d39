#!/bin/bash
# GPU call 14: tests with the serial copy mode, then the end-to-end leg at N = 1 in both modes
mkdir -p gpurun_out
python -m pytest tests/test_parity.py tests/test_abi.py -m gpu -x -q 2>&1 | tail -3
for mode in duplex serial; do
  NTC_COPY_DIRECTIONS=$mode python bench.py --steps 3 --warmup 3 --cpu-budget 0 > gpurun_out/r2_bench_n1_$mode.json 2> gpurun_out/r2_bench_n1_$mode.err
  python - <<PY
import json
s=open("gpurun_out/r2_bench_n1_$mode.json").read(); b=json.loads(s[s.index('{"metric'):]); e=b["e2e"]
print("N=1 $mode value=%.0f e2e=%.0f enc=%.2f ms dec=%.2f ms | %s" % (b["value"], e["value"], e["encode_ms_per_step"], e["decode_ms_per_step"], e["copy_directions"]))
PY
done
