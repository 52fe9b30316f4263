# r1i: bench lines after the host pipeline changes (plan-ahead 2, lazy cigar blocks) + planner kernel profile
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
run() { name=$1; shift; timeout 900 python bench.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"; tail -2 gpurun_out/$name.err; python tools/bench_summary.py < gpurun_out/$name.json; }
run r1i_bench_trna --steps 5 --warmup 3
run r1i_bench_realign --steps 4 --warmup 3 --mode realign --no-cpu-baseline
run r1i_bench_em --steps 3 --warmup 3 --mode em --reads 300000 --no-cpu-baseline
run r1i_bench_long --steps 3 --warmup 3 --workload long --no-cpu-baseline
run r1i_bench_trna3_em --steps 3 --warmup 3 --states 3 --mode em --reads 300000 --no-cpu-baseline
CMD="python bench.py --steps 1 --warmup 1 --reads 200000 --no-cpu-baseline"
ncu --set full --clock-control none --import-source on -k regex:'k_(bands|schedule|regions)' -s 4 -c 4 -o gpurun_out/r1i_planner -f $CMD > gpurun_out/ncu_pl.log 2>&1; echo "ncu planner rc=$?"
