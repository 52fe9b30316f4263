mkdir -p gpurun_out
run() { name=$1; shift; timeout 900 python bench.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"; tail -2 gpurun_out/$name.err; python tools/bench_summary.py < gpurun_out/$name.json; }
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_em.py -m gpu -x -q > gpurun_out/pytest_em.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_em.log
for nb in 4 5 6; do
MAB_EXP_BLOCKS=$nb run em_nb$nb --steps 3 --warmup 3 --mode em --reads 300000 --no-cpu-baseline --no-e2e
MAB_EXP_BLOCKS=$nb run em3_nb$nb --steps 3 --warmup 3 --mode em --states 3 --reads 300000 --no-cpu-baseline --no-e2e
MAB_EXP_BLOCKS=$nb run emlong_nb$nb --steps 3 --warmup 3 --mode em --workload long --no-cpu-baseline --no-e2e
done
