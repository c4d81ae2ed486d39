#!/bin/bash
# N-GPU evidence call: copy-bandwidth probe with all ranks copying, e2e against the chunk size, the bench line
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
$TR --master-port 29533 tools/e2e_probe.py > gpurun_out/r2_e2e_probe_n$N.json 2> gpurun_out/r2_e2e_probe_n$N.err; echo "probe rc=$?"
for cb in 16000000 48000000 192000000; do
  NTC_MAX_CHUNK_BASES=$cb $TR --master-port 29534 bench.py --gpus $N --steps 3 --warmup 3 --cpu-budget 0 --no-extras > gpurun_out/r2_bench_n${N}_chunk$cb.json 2> gpurun_out/r2_bench_n${N}_chunk$cb.err
  python - <<PY
import json
s=open("gpurun_out/r2_bench_n${N}_chunk$cb.json").read(); b=json.loads(s[s.index('{"metric'):]); e=b["e2e"]
print("N=$N chunk_bases=$cb value=%.0f e2e=%.0f enc=%.2f ms dec=%.2f ms h2d=%.1f d2h=%.1f floor=%.1f ms" % (b["value"], e["value"], e["encode_ms_per_step"], e["decode_ms_per_step"], e["pcie_h2d_gbs"], e["pcie_d2h_gbs"], e["pcie_floor_ms_per_step"]))
PY
done
$TR --master-port 29535 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r2_bench_n$N.json 2> gpurun_out/r2_bench_n$N.err; echo "bench rc=$?"
python - <<PY
import json
s=open("gpurun_out/r2_bench_n$N.json").read(); b=json.loads(s[s.index('{"metric'):])
print("N=$N value=%.0f ms=%.2f e2e=%.0f" % (b["value"], b["ms_per_step"], b["e2e"]["value"]))
for k in ("c4","c5"):
    if k in b: print(k, "value=%.0f ms=%.2f l2frac=%.3f e2e=%s" % (b[k]["value"], b[k]["ms_per_step"], b[k]["l2_random"]["frac"], b[k]["e2e"] and round(b[k]["e2e"]["value"])))
PY
