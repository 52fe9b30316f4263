mkdir -p gpurun_out
./tools/pipe_peaks > gpurun_out/pipe_peaks.json; cat gpurun_out/pipe_peaks.json
CMD="python bench.py --steps 1 --warmup 1 --reads 220000 --e2e-steps 1 --no-cpu-baseline"
$CMD > gpurun_out/plain3.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_e2e.csv $CMD > gpurun_out/ncu_l.log 2>&1; echo "ncu rc=$?"
python - <<'PY'
import csv,collections
rows=[r for r in csv.reader(open('gpurun_out/launches_e2e.csv')) if len(r)>10 and r[0].isdigit()]
c=collections.OrderedDict()
for r in rows:
    k=r[4][:60]; c.setdefault(k,[0,0.0]); c[k][0]+=1; c[k][1]+=float(r[-1])/1e6
for k,v in c.items(): print("%4d %9.3f ms  %s"%(v[0],v[1],k))
PY
