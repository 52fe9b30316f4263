mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --mode em --steps 3 --warmup 3 --reads 300000 --no-cpu-baseline > gpurun_out/bench_em.json 2> gpurun_out/bench_em.err; echo "em rc=$?"; tail -3 gpurun_out/bench_em.err; python tools/bench_summary.py < gpurun_out/bench_em.json
