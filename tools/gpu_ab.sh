#!/bin/bash
# pytest under forced affine rounds, then A/B runs of the headline bench under env knobs, optional ncu --set full capture.
mkdir -p gpurun_out
FB_MSM_ROUNDS=3 timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_rounds3.log 2>&1; echo "pytest(rounds=3) rc=$?"; tail -3 gpurun_out/pytest_rounds3.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
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
for spec in $AB; do
  name=${spec%%:*}; envs=${spec#*:}
  run $name $(echo $envs | tr ',' ' ')
done
if [ -n "$NCU_KERNEL" ]; then
env $NCU_ENV timeout 300 python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/plain.log 2>&1 &&
env $NCU_ENV timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$NCU_KERNEL" -s ${NCU_SKIP:-4} -c 2 -f -o gpurun_out/prof_$NCU_TAG python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
fi
