#!/bin/bash
# launch list + optional full capture of one kernel, for the env in $NCU_ENV
mkdir -p gpurun_out
env $NCU_ENV timeout 300 python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/plain.log 2>&1 &&
env $NCU_ENV timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/ncu_list.log 2>&1
echo "ncu list rc=$?"
if [ -n "$NCU_KERNEL" ]; then
env $NCU_ENV timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$NCU_KERNEL" -s ${NCU_SKIP:-4} -c ${NCU_COUNT:-2} -f -o gpurun_out/prof_$NCU_TAG python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
fi
