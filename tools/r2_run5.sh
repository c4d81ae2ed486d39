#!/bin/bash
# GPU call 5: pipeline tests, CLI throughput (mapped + stream input), trace
python -m pytest tests/test_pipeline.py tests/test_container.py tests/test_abi.py -m gpu -x -q 2>&1 | tail -5
NTC_TRACE=1 python tools/cli_bench.py --config C3 --reads 4000000 --tmp /dev/shm > gpurun_out/r2_cli_c3_b.json 2> gpurun_out/r2_cli_c3_b.err; echo rc=$?
cat gpurun_out/r2_cli_c3_b.json; grep -v "^\[ntc trace\] chunk" gpurun_out/r2_cli_c3_b.err | tail -40
NTC_NO_MMAP=1 NTC_TRACE=1 python tools/cli_bench.py --config C3 --reads 4000000 --tmp /dev/shm > gpurun_out/r2_cli_c3_b_stream.json 2> gpurun_out/r2_cli_c3_b_stream.err; echo rc=$?
cat gpurun_out/r2_cli_c3_b_stream.json; grep "encode_file\|fastx" gpurun_out/r2_cli_c3_b_stream.err
