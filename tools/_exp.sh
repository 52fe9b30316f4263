for v in 50 51 40 41 60 61 71; do
  MAB_LA=$v timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "golden_cases or batched_trna" > gpurun_out/r2e_pytest_la$v.log 2>&1; echo "LA=$v pytest rc=$?"
  MAB_LA=$v timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r2e_bench_la$v.json 2>gpurun_out/r2e_bench_la$v.err; python tools/bench_summary.py < gpurun_out/r2e_bench_la$v.json
done
