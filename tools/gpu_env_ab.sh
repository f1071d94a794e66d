#!/bin/bash
# A/B of environment knobs on the headline step (run under gpurun):
#   tools/gpu_env_ab.sh "FB_MSM_REGATHER=1 FB_MSM_PREFETCH=2" "FB_MSM_REGATHER=1" ...
mkdir -p gpurun_out
i=0
for cfg in "$@"; do
  env $cfg python bench.py --steps ${STEPS:-10} --warmup 3 --check ${BENCH_ARGS} > gpurun_out/ab_$i.json 2> gpurun_out/ab_$i.err
  python - "$cfg" $i <<'P'
import json, sys
try:
    d = json.loads(open("gpurun_out/ab_%s.json" % sys.argv[2]).read().strip().splitlines()[-1])
    ph = d.get("phases_ms_per_step") or d.get("phases_ms") or {}
    print("[%s]" % sys.argv[1], round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"], 3), d.get("parity"),
          {k.replace("msm.", ""): round(v, 2) for k, v in ph.items() if v > 0.3})
except Exception as e:
    print("[%s] failed: %s" % (sys.argv[1], e))
P
  i=$((i+1))
done
