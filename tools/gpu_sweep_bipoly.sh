#!/bin/bash
# Window / rounds sweep of the bivariate workload (2^10 rows x 2^12) on one GPU (run under gpurun).
mkdir -p gpurun_out
for c in ${WINDOWS:-10 11 12 13}; do
  for r in ${ROUNDS:-3}; do
    FB_BENCH_WINDOW=$c FB_MSM_ROUNDS=$r python bench.py --workload bipoly --steps 3 --warmup 3 > gpurun_out/bsweep_c${c}_r${r}.json 2> gpurun_out/bsweep_c${c}_r${r}.err
    python - "$c" "$r" <<'P'
import json, sys
c, r = sys.argv[1:3]
try:
    d = json.loads(open(f"gpurun_out/bsweep_c{c}_r{r}.json").read().strip().splitlines()[-1])
    ph = d.get("phases_ms", {})
    top = sorted(ph.items(), key=lambda kv: -kv[1])[:6]
    print("c", c, "rounds", r, "ms", round(d["ms_per_step"], 2), "e2e", round(d["e2e"]["value"], 2), [(k, round(v, 1)) for k, v in top])
except Exception as e:
    print("c", c, "rounds", r, "failed", e)
P
  done
done
