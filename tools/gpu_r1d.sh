# r1d: realign tail on the GPU + auto chunk sizing
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/pytest_gpu.log
run() { name=$1; shift; timeout 600 python bench.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"; tail -2 gpurun_out/$name.err; python tools/bench_summary.py < gpurun_out/$name.json; }
run long --steps 3 --warmup 3 --workload long --no-cpu-baseline
run long_em --steps 2 --warmup 3 --workload long --mode em --no-cpu-baseline --no-e2e
run long200 --steps 2 --warmup 3 --workload long --expansion 200 --reads 600 --no-cpu-baseline --no-e2e
