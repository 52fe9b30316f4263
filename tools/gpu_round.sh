mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_trna.json 2> gpurun_out/bench_trna.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_trna.err; python tools/bench_summary.py < gpurun_out/bench_trna.json; cat gpurun_out/bench_trna.json
