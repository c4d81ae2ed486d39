#!/bin/bash
# N-GPU call: end-to-end leg in both copy-direction modes (same box, back to back), then the default bench line
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
for mode in duplex serial duplex serial; do
  NTC_COPY_DIRECTIONS=$mode $TR --master-port 29534 bench.py --gpus $N --steps 3 --warmup 3 --cpu-budget 0 --no-extras > gpurun_out/r2_bench_n${N}_$mode.json 2> gpurun_out/r2_bench_n${N}_$mode.err
  python - <<PY
import json
s=open("gpurun_out/r2_bench_n${N}_$mode.json").read(); b=json.loads(s[s.index('{"metric'):]); e=b["e2e"]
print("N=$N $mode value=%.0f e2e=%.0f enc=%.2f ms dec=%.2f ms h2d=%.1f d2h=%.1f floor=%.1f ms" % (b["value"], e["value"], e["encode_ms_per_step"], e["decode_ms_per_step"], e["pcie_h2d_gbs"], e["pcie_d2h_gbs"], e["pcie_floor_ms_per_step"]))
PY
done
