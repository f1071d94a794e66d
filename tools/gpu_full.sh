#!/bin/bash
# full check with library defaults: GPU tests, smoke, headline bench (own + reference arm), configs 3-5
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"; tail -2 gpurun_out/bench_n1.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
for w in ${WORKLOADS:-ntt msm bipoly}; do
  timeout 900 python bench.py --workload $w --steps ${STEPS:-3} --warmup 3 --check > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "$w rc=$?"; tail -3 gpurun_out/bench_$w.err
done
python - <<'PY'
import json
for w in ("n1","ntt","msm","bipoly"):
    try:
        d=json.loads(open(f"gpurun_out/bench_{w}.json").read().strip().splitlines()[-1])
        print(w, d["metric"], round(d["value"],3), d["unit"], "e2e", d.get("e2e") and round(d["e2e"]["value"],3), d.get("parity"), d.get("phases_ms_per_step") or d.get("phases_ms"))
    except Exception as e: print(w, "failed", e)
PY
