mkdir -p gpurun_out
export FB_BENCH_STEP=separate
CMD="python bench.py --steps 2 --warmup 1 --no-cpu --no-extra"
timeout 900 ncu --clock-control none -k regex:"k_round_pass[12]" -s 12 -c 6 --metrics sm__cycles_active.min,sm__cycles_active.max,sm__cycles_active.avg,sm__cycles_elapsed.max,sm__inst_executed.min,sm__inst_executed.max,sm__inst_executed.avg,sm__warps_active.min,sm__warps_active.max,sm__warps_active.avg,smsp__cycles_active.min,smsp__cycles_active.max,smsp__cycles_active.avg,smsp__inst_executed.min,smsp__inst_executed.max,smsp__inst_executed.avg,gpu__time_duration.sum --csv --log-file gpurun_out/balance.csv $CMD > gpurun_out/balance.log 2>&1
echo rc=$?
python - <<'PY'
import csv
rows=list(csv.reader(open("gpurun_out/balance.csv")))
hi=[i for i,r in enumerate(rows) if r and r[0]=="ID"][0]
h=rows[hi]; ki=h.index("Kernel Name"); mi=h.index("Metric Name"); vi=h.index("Metric Value")
cur=None
for r in rows[hi+1:]:
    if len(r)<=vi: continue
    key=(r[0], r[ki][:40])
    if key!=cur: print("---",key); cur=key
    print("   %-40s %s"%(r[mi], r[vi]))
PY
