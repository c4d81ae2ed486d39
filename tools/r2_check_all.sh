#!/bin/bash
# whole GPU suite + smoke + default bench line + reference arm (short) -- run after changes that touch shared code
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/gpu_tests.log 2>&1; echo "gpu tests rc=$?"; tail -3 gpurun_out/gpu_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_c3.json'))
print({k:d[k] for k in ("value","ms_per_step","gpu_launches")}, "e2e", d["e2e"]["value"], "frac", d["roofline"]["frac"])
print("index_build", d.get("index_build"))
print("cpu", d.get("cpu_baseline",{}) and d["cpu_baseline"]["value"], "setup", d["setup_s"])
PY
