#!/bin/bash
# A/B runs of the headline bench under env knobs, then one ncu --set full capture.
mkdir -p gpurun_out
run() { # name, env...
  name=$1; shift
  env "$@" timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/ab_$name.json 2> gpurun_out/ab_$name.err || tail -5 gpurun_out/ab_$name.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/ab_$name.json").read().strip().splitlines()[-1])
    print("$name", "step_ms", round(d["value"],3), "e2e", round(d["e2e"]["value"],3), {k:v for k,v in d["phases_ms_per_step"].items()})
except Exception as e: print("$name failed", e)
PY
}
run c20_occ3 FB_BENCH_WINDOW=20 FB_ACC_MINB=3
run c20_occ4 FB_BENCH_WINDOW=20 FB_ACC_MINB=4
run c17_occ4 FB_BENCH_WINDOW=17 FB_ACC_MINB=4
if [ -n "$NCU_KERNEL" ]; then
FB_BENCH_WINDOW=20 timeout 300 python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/plain.log 2>&1 &&
FB_BENCH_WINDOW=20 timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$NCU_KERNEL" -s 4 -c 2 -f -o gpurun_out/prof_$NCU_TAG python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
fi
