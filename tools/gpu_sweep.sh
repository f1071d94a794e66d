#!/bin/bash
# Window / affine-round sweep of the headline step on one GPU (run under gpurun).
# Prints one line per setting: window bits, rounds, device ms, e2e ms, parity.
mkdir -p gpurun_out
for c in ${WINDOWS:-16 17 18 19 20}; do
  for r in ${ROUNDS:-3}; do
    FB_BENCH_WINDOW=$c FB_MSM_ROUNDS=$r python bench.py --steps 10 --warmup 3 --check > gpurun_out/sweep_c${c}_r${r}.json 2> gpurun_out/sweep_c${c}_r${r}.err
    python - "$c" "$r" <<'P'
import json, sys
c, r = sys.argv[1:3]
try:
    d = json.loads(open(f"gpurun_out/sweep_c{c}_r{r}.json").read().strip().splitlines()[-1])
    print("c", c, "rounds", r, "ms", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"], 3), d.get("parity"))
except Exception as e:
    print("c", c, "rounds", r, "failed", e)
P
  done
done
