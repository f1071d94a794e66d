#!/bin/bash
# multi-GPU runs on one box: bash tools/gpu_scale.sh N  (N ranks under torchrun; needs gpurun --gpus N)
N=${1:-2}
mkdir -p gpurun_out
run() { # tag, args...
  tag=$1; shift
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N "$@" > gpurun_out/scale_${tag}_n$N.json 2> gpurun_out/scale_${tag}_n$N.err
  echo "$tag n=$N rc=$?"; grep -v "^\*\*\*\|OMP_NUM\|^$\|NCCL version" gpurun_out/scale_${tag}_n$N.err | tail -3
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/scale_${tag}_n$N.json").read().strip().splitlines()[-1])
    print("  ", d["metric"], round(d["value"],3), d["unit"], "e2e", d.get("e2e") and round(d["e2e"]["value"],3), d.get("parity"))
except Exception as e: print("  failed", e)
PY
}
run kzg --steps 10 --warmup 3
for w in ${WORKLOADS-msm bipoly ntt}; do run $w --workload $w --steps 3 --warmup 3; done
