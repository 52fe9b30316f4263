# quick check: GPU parity tests + default bench line
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_trna.json 2> gpurun_out/bench_trna.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_trna.err; python tools/bench_summary.py < gpurun_out/bench_trna.json
timeout 600 python bench.py --steps 3 --warmup 3 --mode em --reads 300000 --no-cpu-baseline > gpurun_out/bench_em.json 2> gpurun_out/bench_em.err; echo "em rc=$?"; python tools/bench_summary.py < gpurun_out/bench_em.json
