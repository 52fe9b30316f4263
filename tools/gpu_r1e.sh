# r1e: warp-cooperative posterior kernel, realign mode in bench
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/pytest_gpu.log
run() { name=$1; shift; timeout 900 python bench.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"; tail -2 gpurun_out/$name.err; python tools/bench_summary.py < gpurun_out/$name.json; }
run trna --steps 3 --warmup 3 --no-cpu-baseline
run trna_realign --steps 3 --warmup 3 --mode realign --no-cpu-baseline
run long --steps 3 --warmup 3 --workload long --no-cpu-baseline --no-e2e
run long200 --steps 2 --warmup 3 --workload long --expansion 200 --reads 600 --no-cpu-baseline --no-e2e
run long_realign --steps 2 --warmup 3 --workload long --mode realign --no-cpu-baseline
timeout 600 python bench.py --impl reference --mode realign --steps 1 --warmup 1 > gpurun_out/ref_realign.json 2> gpurun_out/ref_realign.err; echo "ref realign rc=$?"; cat gpurun_out/ref_realign.json | cut -c1-400
