#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_pipeline.py -x -q -m gpu > gpurun_out/pipeline_tests.log 2>&1; echo "tests rc=$?"; tail -2 gpurun_out/pipeline_tests.log
for bb in 2 4 8; do
NTC_TRACE=1 NTC_PIPELINE_BATCH_BLOCKS=$bb timeout 600 python tools/cli_bench.py --config C3 --reads 4000000 --tmp /dev/shm > gpurun_out/cli_c3_inflight_batch$bb.json 2> gpurun_out/cli_$bb.err
python - <<PY
import json
s=open("gpurun_out/cli_c3_inflight_batch$bb.json").read(); d=json.loads(s[s.index('{'):])
print("batch_blocks=$bb encode %.3f s (%.2f Gbases/s) decode %.3f s sha %s" % (d["encode_pipeline_s"], d["encode_gbases_per_s"], d["decode_pipeline_s"], d["encoded_sha256"][:12]))
print("  ", d["trace"]["encode"][-1][:230])
PY
done
