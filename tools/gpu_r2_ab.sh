#!/bin/bash
# A/B runs of the headline bench (and optionally the msm / bipoly workloads) under env knobs.
#   AB="name:ENV=1,ENV2=2 name2:..."   WAB="msm:name:ENV=.. bipoly:name:ENV=.."   PYTEST_K="expr" to run a test subset first
mkdir -p gpurun_out
if [ -n "$PYTEST_K" ]; then
  timeout 1200 python -m pytest tests -m gpu -x -q -k "$PYTEST_K" > gpurun_out/pytest_k.log 2>&1; echo "pytest -k rc=$?"; tail -5 gpurun_out/pytest_k.log
fi
for spec in $AB; do
  name=${spec%%:*}; envs=${spec#*:}
  env $(echo $envs | tr ',' ' ') timeout 300 python bench.py --steps ${STEPS:-10} --warmup 3 --no-cpu --no-extra > gpurun_out/ab_$name.json 2> gpurun_out/ab_$name.err || tail -5 gpurun_out/ab_$name.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/ab_$name.json").read().strip().splitlines()[-1])
    print("$name", d["extra"]["step_variants_ms"], "e2e", {k:round(v,2) for k,v in d["e2e"]["all_ms"].items()})
    print("    phases", d["phases_ms_per_step"])
except Exception as e: print("$name failed", e)
PY
done
for spec in $WAB; do
  w=${spec%%:*}; rest=${spec#*:}; name=${rest%%:*}; envs=${rest#*:}
  env $(echo $envs | tr ',' ' ') timeout 600 python bench.py --workload $w --steps ${WSTEPS:-3} --warmup 3 > gpurun_out/wab_${w}_$name.json 2> gpurun_out/wab_${w}_$name.err || tail -5 gpurun_out/wab_${w}_$name.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/wab_${w}_$name.json").read().strip().splitlines()[-1])
    print("$w $name", d["metric"], round(d["value"],3), d["unit"], "e2e", round(d["e2e"]["value"],3))
    print("    phases", d.get("phases_ms"))
except Exception as e: print("$w $name failed", e)
PY
done
