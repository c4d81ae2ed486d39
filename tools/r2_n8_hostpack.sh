#!/bin/bash
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
for t in 0 4; do
NTC_HOST_THREADS=$t $TR --master-port 2953$t bench.py --gpus $N --steps 3 --warmup 3 --cpu-budget 0 --no-extras > gpurun_out/bench_c3_n${N}_hostpack$t.json 2> gpurun_out/bench_c3_n${N}_hostpack$t.err
python - <<PY
import json
s=open("gpurun_out/bench_c3_n${N}_hostpack$t.json").read(); d=json.loads(s[s.index('{"metric'):]); e=d["e2e"]
print("N=$N host_threads=$t value=%.0f e2e=%.0f enc=%.2f ms dec=%.2f ms h2d_gbs=%.1f d2h_gbs=%.1f binding=%s" % (d["value"], e["value"], e["encode_ms_per_step"], e["decode_ms_per_step"], e["pcie_h2d_gbs"], e["pcie_d2h_gbs"], d["host_binding"]))
PY
done
