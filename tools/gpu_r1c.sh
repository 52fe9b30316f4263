# r1c: warp-group wavefront kernels for wide bands -- parity + long-read sweeps
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
for g in 1 2 4 8; do
  MAB_W5_GROUP=$g timeout 600 python bench.py --steps 3 --warmup 3 --workload long --no-cpu-baseline --no-e2e > gpurun_out/long_g$g.json 2> gpurun_out/long_g$g.err; echo "long g=$g rc=$?"; python tools/bench_summary.py < gpurun_out/long_g$g.json
done
timeout 600 python bench.py --steps 3 --warmup 3 --workload long --no-cpu-baseline > gpurun_out/bench_long.json 2> gpurun_out/bench_long.err; echo "long rc=$?"; python tools/bench_summary.py < gpurun_out/bench_long.json
timeout 600 python bench.py --steps 2 --warmup 3 --workload long --mode em --no-cpu-baseline --no-e2e > gpurun_out/bench_long_em.json 2> gpurun_out/bench_long_em.err; echo "long em rc=$?"; python tools/bench_summary.py < gpurun_out/bench_long_em.json
timeout 600 python bench.py --steps 2 --warmup 3 --workload long --expansion 200 --reads 600 --no-cpu-baseline --no-e2e > gpurun_out/bench_long200.json 2> gpurun_out/bench_long200.err; echo "long200 rc=$?"; python tools/bench_summary.py < gpurun_out/bench_long200.json
MAB_W5_GROUP=1 timeout 600 python bench.py --steps 2 --warmup 3 --workload long --expansion 200 --reads 600 --no-cpu-baseline --no-e2e > gpurun_out/bench_long200_g1.json 2> gpurun_out/bench_long200_g1.err; echo "long200 g1 rc=$?"; python tools/bench_summary.py < gpurun_out/bench_long200_g1.json
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_trna.json 2> gpurun_out/bench_trna.err; echo "trna rc=$?"; python tools/bench_summary.py < gpurun_out/bench_trna.json
