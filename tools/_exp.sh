timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2i_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2i_pytest.log
for v in 5 4 6; do
  MAB_POST_MINB=$v timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-em-leg --no-strong-leg > gpurun_out/r2i_bench_minb$v.json 2>gpurun_out/r2i_bench_minb$v.err; echo "MINB=$v"; python tools/bench_summary.py < gpurun_out/r2i_bench_minb$v.json
done
