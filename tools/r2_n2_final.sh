#!/bin/bash
# N-GPU bench line exactly as the driver launches it (c4 object at N > 1, c5 at N = 8), plus the reference arm under torchrun
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
$TR --master-port 29535 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_c3_n${N}_r2b.json 2> gpurun_out/bench_c3_n${N}_r2b.err; echo "bench rc=$?"
python - <<PY
import json
s=open("gpurun_out/bench_c3_n${N}_r2b.json").read(); b=json.loads(s[s.index('{"metric'):])
print("N=$N value=%.0f ms=%.2f e2e=%.0f index_build=%s" % (b["value"], b["ms_per_step"], b["e2e"]["value"], b.get("index_build")))
for k in ("c4","c5"):
    if k in b: print(k, "value=%.0f ms=%.2f l2frac=%.3f e2e=%s" % (b[k]["value"], b[k]["ms_per_step"], b[k]["l2_random"]["frac"], b[k]["e2e"] and round(b[k]["e2e"]["value"])))
PY
tail -3 gpurun_out/bench_c3_n${N}_r2b.err
