# r1c: per-kernel warp-group choice + segment work items in k_expect
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
run() { name=$1; shift; timeout 600 python bench.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"; python tools/bench_summary.py < gpurun_out/$name.json; }
run long --steps 3 --warmup 3 --workload long --no-cpu-baseline --no-e2e
run long_em --steps 2 --warmup 3 --workload long --mode em --no-cpu-baseline --no-e2e
run long200 --steps 2 --warmup 3 --workload long --expansion 200 --reads 600 --no-cpu-baseline --no-e2e
run long500_em --steps 2 --warmup 2 --workload long --expansion 500 --reads 100 --mode em --no-cpu-baseline --no-e2e
run trna_em --steps 3 --warmup 3 --mode em --reads 300000 --no-cpu-baseline --no-e2e
