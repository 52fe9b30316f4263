# r1i: record-based chain kernel
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_gpu.log
run() { name=$1; shift; timeout 900 python bench.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"; tail -2 gpurun_out/$name.err; python tools/bench_summary.py < gpurun_out/$name.json; }
run trna_realign --steps 3 --warmup 3 --mode realign --no-cpu-baseline
run long_realign --steps 2 --warmup 3 --workload long --mode realign --no-cpu-baseline --no-e2e
