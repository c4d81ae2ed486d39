#!/bin/bash
# multi-GPU evidence on one box: tools/scale_run.sh TAG N [config ...]
TAG=$1; N=$2; shift 2
mkdir -p gpurun_out
for cfg in "$@"; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 3 --warmup 3 --cpu-budget 0 --config $cfg > gpurun_out/bench_${cfg,,}_n${N}_$TAG.json 2> gpurun_out/bench_${cfg,,}_n${N}_$TAG.err
  echo "$cfg N=$N rc=$?"; head -c 400 gpurun_out/bench_${cfg,,}_n${N}_$TAG.json; echo
done
