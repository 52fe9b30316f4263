mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "device_planner or host_and_device or golden" > gpurun_out/pytest_plan.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_plan.log
CMD="python bench.py --steps 1 --warmup 1 --reads 110000 --e2e-steps 1 --no-cpu-baseline"
$CMD > gpurun_out/plain4.log 2>&1 && ncu --metrics gpu__time_duration.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__inst_executed.sum --clock-control none -k regex:'k_(regions|bands|schedule|symbols)' -s 6 -c 6 --csv --log-file gpurun_out/plan_k.csv $CMD > gpurun_out/ncu_p.log 2>&1; echo "ncu rc=$?"
python - <<'PY'
import csv
for r in csv.reader(open('gpurun_out/plan_k.csv')):
    if len(r)>10 and r[0].isdigit(): print(r[4][:50], r[-3], r[-1])
PY
