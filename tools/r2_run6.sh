#!/bin/bash
# GPU call 6: pipeline tests, CLI throughput at batch 4 / 8 blocks and gzip level 1 (information)
python -m pytest tests/test_pipeline.py tests/test_container.py tests/test_abi.py -m gpu -x -q 2>&1 | tail -5
for v in "a:" "b:NTC_PIPELINE_BATCH_BLOCKS=8" "c:NTC_PIPELINE_BATCH_BLOCKS=2" "d:NTC_GZIP_LEVEL=1" ; do
  tag=${v%%:*}; envs=${v#*:}
  env $envs NTC_TRACE=1 python tools/cli_bench.py --config C3 --reads 4000000 --tmp /dev/shm > gpurun_out/r2_cli_c3_c_$tag.json 2> gpurun_out/r2_cli_c3_c_$tag.err; echo rc=$?
  python - <<PY
import json
d=json.load(open("gpurun_out/r2_cli_c3_c_$tag.json"))
print("$envs", d["encode_pipeline_s"], d["encode_setup_s"], d["decode_pipeline_s"], d["decode_setup_s"], d["encoded_bytes"], d["encoded_sha256"][:12], d["round_trip_ok"])
for l in d["trace"]["encode"]+d["trace"]["decode"]: print("   ", l)
PY
done
