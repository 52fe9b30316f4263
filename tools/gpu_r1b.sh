# Round-1 state capture: GPU parity tests, bench lines, ncu launch list, one ncu --set full of the DP kernels.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt
nproc >> gpurun_out/gpu.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_trna.json 2> gpurun_out/bench_trna.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_trna.err; cat gpurun_out/bench_trna.json
timeout 600 python bench.py --steps 3 --warmup 3 --mode em --reads 300000 --no-cpu-baseline > gpurun_out/bench_em.json 2> gpurun_out/bench_em.err; echo "em rc=$?"; cat gpurun_out/bench_em.json
timeout 600 python bench.py --steps 3 --warmup 3 --workload long --no-cpu-baseline > gpurun_out/bench_long.json 2> gpurun_out/bench_long.err; echo "long rc=$?"; cat gpurun_out/bench_long.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cat gpurun_out/bench_ref.json
CMD="python bench.py --steps 1 --warmup 1 --reads 200000 --no-e2e --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_l.log 2>&1; echo "ncu list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'k_(forward|backward)5' -s 2 -c 2 -o gpurun_out/dp_full -f $CMD > gpurun_out/ncu_f.log 2>&1; echo "ncu full rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'k_(posterior_diag|emit_diag|totals)' -s 3 -c 3 -o gpurun_out/post_full -f $CMD > gpurun_out/ncu_p.log 2>&1; echo "ncu post rc=$?"
ls -la gpurun_out
