# r1j: sync-free, software-pipelined host paths (posterior + realign)
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_gpu.log
run() { name=$1; shift; timeout 900 python bench.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"; tail -2 gpurun_out/$name.err; python tools/bench_summary.py < gpurun_out/$name.json; }
run trna --steps 3 --warmup 3 --no-cpu-baseline
run trna_realign --steps 3 --warmup 3 --mode realign --no-cpu-baseline
run long --steps 3 --warmup 3 --workload long --no-cpu-baseline
