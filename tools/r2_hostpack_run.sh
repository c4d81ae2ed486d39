#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_parity.py tests/test_sharding.py tests/test_pipeline.py -x -q -m gpu > gpurun_out/hostpack_tests.log 2>&1; echo "tests rc=$?"; tail -2 gpurun_out/hostpack_tests.log
for t in 0 4 8 16; do
NTC_HOST_THREADS=$t timeout 600 python bench.py --cpu-budget 0 --steps 3 > gpurun_out/bench_c3_hostpack$t.json 2> gpurun_out/bench_c3_hostpack$t.err
python - <<PY
import json
d=json.load(open("gpurun_out/bench_c3_hostpack$t.json")); e=d["e2e"]
print("host_threads=$t value=%.0f e2e=%.0f enc=%.2f ms dec=%.2f ms h2d=%d floor=%.1f" % (d["value"], e["value"], e["encode_ms_per_step"], e["decode_ms_per_step"], e["h2d_bytes_per_step"], e["pcie_floor_ms_per_step"]))
PY
done
