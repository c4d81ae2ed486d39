#!/bin/bash
mkdir -p gpurun_out
for t in 16 8 4 0; do
NTC_HOST_THREADS=$t timeout 600 python bench.py --cpu-budget 0 --steps 3 > gpurun_out/bench_c3_hostpack$t.json 2> gpurun_out/bench_c3_hostpack$t.err
python - <<PY
import json
d=json.load(open("gpurun_out/bench_c3_hostpack$t.json")); e=d["e2e"]
print("host_threads=$t value=%.0f e2e=%.0f enc=%.2f ms dec=%.2f ms h2d=%d floor=%.1f" % (d["value"], e["value"], e["encode_ms_per_step"], e["decode_ms_per_step"], e["h2d_bytes_per_step"], e["pcie_floor_ms_per_step"]))
PY
done
NTC_HOST_THREADS=16 NTC_MAX_CHUNK_BASES=200000000 timeout 600 python bench.py --cpu-budget 0 --steps 3 > gpurun_out/bench_c3_hostpack16_big.json 2>/dev/null
python - <<PY
import json
d=json.load(open("gpurun_out/bench_c3_hostpack16_big.json")); e=d["e2e"]
print("host_threads=16 chunk=200M value=%.0f e2e=%.0f enc=%.2f ms dec=%.2f ms" % (d["value"], e["value"], e["encode_ms_per_step"], e["decode_ms_per_step"]))
PY
