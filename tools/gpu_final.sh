# round-1 final state capture: GPU suite, smoke, bench lines (r1h), ncu launch lists, one ncu --set full of the dominant kernels
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt; nproc >> gpurun_out/gpu.txt
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
run() { name=$1; shift; timeout 900 python bench.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"; tail -2 gpurun_out/$name.err; python tools/bench_summary.py < gpurun_out/$name.json; }
run r1h_bench_trna --steps 5 --warmup 3
run r1h_bench_realign --steps 3 --warmup 3 --mode realign
run r1h_bench_em --steps 3 --warmup 3 --mode em --reads 300000 --no-cpu-baseline
run r1h_bench_long --steps 3 --warmup 3 --workload long --no-cpu-baseline
run r1h_bench_long_em --steps 3 --warmup 3 --workload long --mode em --no-cpu-baseline
run r1h_bench_long200 --steps 2 --warmup 3 --workload long --expansion 200 --reads 600 --no-cpu-baseline --no-e2e
run r1h_bench_trna3 --steps 3 --warmup 3 --states 3 --no-cpu-baseline
run r1h_bench_trna3_em --steps 3 --warmup 3 --states 3 --mode em --reads 300000 --no-cpu-baseline
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r1h_bench_ref.json 2> gpurun_out/r1h_bench_ref.err; echo "ref rc=$?"
timeout 600 python bench.py --impl reference --mode realign --steps 1 --warmup 1 > gpurun_out/r1h_bench_ref_realign.json 2>> gpurun_out/r1h_bench_ref.err; echo "ref realign rc=$?"
CMD="python bench.py --steps 1 --warmup 1 --reads 200000 --mode realign --no-e2e --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1h_launches_trna_realign.csv $CMD > gpurun_out/ncu_l.log 2>&1; echo "ncu list rc=$?"; python tools/bench_summary.py < gpurun_out/plain.log
ncu --set full --clock-control none --import-source on -k regex:'k_(forward|backward)$' -s 2 -c 2 -o gpurun_out/r1h_dp_trna -f $CMD > gpurun_out/ncu_f.log 2>&1; echo "ncu full trna rc=$?"
CMD3="python bench.py --steps 1 --warmup 1 --reads 200000 --states 3 --no-e2e --no-cpu-baseline"
$CMD3 > gpurun_out/plain3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'k_(forward|backward)' -s 2 -c 2 -o gpurun_out/r1h_dp_trna3 -f $CMD3 > gpurun_out/ncu_3.log 2>&1; echo "ncu full 3-state rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'k_(totals|posterior_diag|rl_chain|rl_weights)' -s 8 -c 4 -o gpurun_out/r1h_post_trna -f $CMD > gpurun_out/ncu_p.log 2>&1; echo "ncu post rc=$?"
