#!/bin/bash
# end-of-round evidence on one B200 (tag r2f): GPU tests, smoke, bench (both arms), the other single-GPU configs, ncu launch list of the
# bench command, ncu --set full of the dominant kernels (each ncu pass after the same command ran clean without it)
TAG=${1:-r2f}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -4 | tee gpurun_out/gpu_tests_$TAG.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1 | tee -a gpurun_out/gpu_tests_$TAG.log
python bench.py > gpurun_out/bench_c3_$TAG.json 2> gpurun_out/bench_c3_$TAG.err; echo "bench rc=$?"
python bench.py --impl reference > gpurun_out/bench_c3_reference_arm_$TAG.json 2> gpurun_out/bench_ref_$TAG.err; echo "ref rc=$?"
for cfg in C2 C4 C5; do python bench.py --config $cfg --cpu-budget 0 --steps 3 > gpurun_out/bench_${cfg,,}_$TAG.json 2> gpurun_out/bench_${cfg,,}_$TAG.err; echo "$cfg rc=$?"; done
python bench.py --steps 2 --warmup 3 --cpu-budget 0 --no-e2e > gpurun_out/bench_plain_$TAG.json 2>/dev/null && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/launches_$TAG.csv python bench.py --steps 2 --warmup 3 --cpu-budget 0 --no-e2e > gpurun_out/ncu_launch_$TAG.log 2>&1
echo "ncu launch list rc=$?"
python bench.py --steps 1 --warmup 3 --cpu-budget 0 --no-e2e > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'ms_kernel|walk_kernel|emit_kernel|pack2bit_kernel' -s 16 -c 4 -f -o gpurun_out/prof_$TAG python bench.py --steps 1 --warmup 3 --cpu-budget 0 --no-e2e > gpurun_out/ncu_full_$TAG.log 2>&1
echo "ncu full rc=$?"; ls -la gpurun_out/prof_$TAG.ncu-rep
python - <<PY
import json
for f in ("bench_c3_$TAG","bench_c3_reference_arm_$TAG","bench_c2_$TAG","bench_c4_$TAG","bench_c5_$TAG"):
    try:
        s=open("gpurun_out/%s.json"%f).read(); b=json.loads(s[s.index('{"'):])
        print(f, "value=%.0f"%b["value"], "ms=%.2f"%b.get("ms_per_step",0), "e2e=%.0f"%(b["e2e"]["value"] if b.get("e2e") else 0), "frac=%s"%(b.get("roofline") or {}).get("frac"))
    except Exception as e: print(f, "ERR", e)
PY
