#!/bin/bash
# GPU tests + window sweep of the headline bench + launch list (ncu)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
for c in ${SWEEP:-0 15 17 18 20 21}; do
  FB_BENCH_WINDOW=$c timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/sweep_c$c.json 2> gpurun_out/sweep_c$c.err || tail -5 gpurun_out/sweep_c$c.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/sweep_c$c.json").read().strip().splitlines()[-1])
    print("c=$c", "step_ms", round(d["value"],3), "e2e", round(d["e2e"]["value"],3), {k:v for k,v in d["phases_ms_per_step"].items()})
except Exception as e: print("c=$c failed", e)
PY
done
if [ -n "$NCU_LIST" ]; then
FB_BENCH_WINDOW=$NCU_LIST timeout 300 python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/plain.log 2>&1 &&
FB_BENCH_WINDOW=$NCU_LIST timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/ncu_list.log 2>&1
echo "ncu list rc=$?"
fi
