mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu.log
run() { name=$1; shift; timeout 900 python bench.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"; tail -2 gpurun_out/$name.err; python tools/bench_summary.py < gpurun_out/$name.json; }
run trna3 --steps 3 --warmup 3 --states 3 --no-cpu-baseline --no-e2e
run trna3_em --steps 3 --warmup 3 --states 3 --mode em --reads 300000 --no-cpu-baseline --no-e2e
run long3 --steps 3 --warmup 3 --states 3 --workload long --no-cpu-baseline --no-e2e
