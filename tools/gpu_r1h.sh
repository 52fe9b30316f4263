# r1h: full GPU suite after the sentinel-warp fix, launch lists of the fixed state, default bench line
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_gpu.log
CMD="python bench.py --steps 1 --warmup 1 --reads 200000 --mode realign --no-e2e --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1f_launches_trna_realign.csv $CMD > gpurun_out/ncu_l.log 2>&1; echo "ncu list rc=$?"; python tools/bench_summary.py < gpurun_out/plain.log
CMDL="python bench.py --steps 1 --warmup 1 --workload long --reads 1000 --no-e2e --no-cpu-baseline"
$CMDL > gpurun_out/plainl.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1f_launches_long.csv $CMDL > gpurun_out/ncu_ll.log 2>&1; echo "ncu list long rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'k_(posterior_diag|rl_chain|totals)' -s 6 -c 3 -o gpurun_out/r1f_post_trna -f $CMD > gpurun_out/ncu_p.log 2>&1; echo "ncu post rc=$?"
