#!/bin/bash
# GPU tests, then the other BASELINE workloads (configs 3-5) on one GPU
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
for w in ${WORKLOADS:-ntt msm bipoly}; do
  timeout 900 python bench.py --workload $w --steps ${STEPS:-3} --warmup 3 --check > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "$w rc=$?"; tail -3 gpurun_out/bench_$w.err
  head -c 2500 gpurun_out/bench_$w.json; echo
done
