#!/bin/bash
# round 2: launch list of the headline step + one `ncu --set full` capture of the MSM kernels of one MSM (and optionally of
# the NTT).  The plain run (same command line, no ncu) goes first, as the profiling recipe asks.
#   NCU_KERNEL (regex), NCU_SKIP, NCU_COUNT, NCU_TAG; NCU_NTT=1 adds a capture of k_ntt_pass from the NTT workload
mkdir -p gpurun_out
export FB_BENCH_STEP=${FB_BENCH_STEP:-separate}
CMD="python bench.py --steps 2 --warmup 1 --no-cpu --no-extra"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "ncu list rc=$?"
if [ -n "$NCU_KERNEL" ]; then
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"$NCU_KERNEL" -s ${NCU_SKIP:-34} -c ${NCU_COUNT:-20} -f -o gpurun_out/prof_${NCU_TAG:-r2_msm} $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
fi
if [ -n "$NCU_NTT" ]; then
NCMD="python bench.py --workload ntt --steps 2 --warmup 1"
timeout 300 $NCMD > gpurun_out/plain_ntt.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_ntt_pass -s ${NCU_NTT_SKIP:-60} -c 2 -f -o gpurun_out/prof_${NCU_TAG:-r2}_ntt $NCMD > gpurun_out/ncu_ntt.log 2>&1
echo "ncu ntt rc=$?"
fi
# summarise on the box (ncu is here too) and drop reports that would push gpurun_out/ over the 64 MiB that travel back
python tools/ncu_summary.py launches gpurun_out/launches.csv > gpurun_out/${NCU_TAG:-r2}_launches.txt 2>/dev/null
for f in gpurun_out/*.ncu-rep; do
  [ -f "$f" ] || continue
  python tools/ncu_summary.py full $f > ${f%.ncu-rep}.txt 2>/dev/null
  python tools/ncu_summary.py source $f > ${f%.ncu-rep}_source.txt 2>/dev/null
  sz=$(stat -c %s $f); if [ $sz -gt 25000000 ]; then rm -f $f; echo "dropped $f ($sz bytes; summary kept)"; fi
done
ls -la gpurun_out/
