#!/bin/bash
# round 2: GPU tests, smoke, headline bench (own + reference arm); optional workloads via WORKLOADS="msm ntt bipoly"
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv,noheader
timeout ${PYTEST_TIMEOUT:-2400} python -m pytest tests -m gpu -x -q ${PYTEST_ARGS} --durations=15 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 900 python bench.py --steps ${STEPS:-10} --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"; tail -5 gpurun_out/bench_n1.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
for w in ${WORKLOADS}; do
  timeout 900 python bench.py --workload $w --steps ${WSTEPS:-3} --warmup 3 --check > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "$w rc=$?"; tail -3 gpurun_out/bench_$w.err
done
python - <<'PY'
import json
for w in ("n1","ref","ntt","msm","bipoly"):
    try:
        d=json.loads(open(f"gpurun_out/bench_{w}.json").read().strip().splitlines()[-1])
        print(w, d["metric"], round(d["value"],3), d["unit"], "e2e", d.get("e2e") and d["e2e"].get("all_ms", d["e2e"]["value"]), d.get("parity"))
        print("   phases", d.get("phases_ms_per_step") or d.get("phases_ms"))
        if "extra" in d: print("   extra", json.dumps(d["extra"])[:3000])
        if "roofline" in d: print("   roofline frac", d["roofline"].get("frac"), "frac_real", d["roofline"].get("frac_real"), d["roofline"].get("dominant_kernel"))
    except Exception as e: print(w, "failed", e)
PY
