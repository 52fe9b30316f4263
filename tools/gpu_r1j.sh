# final tree check: GPU suite, smoke, the driver's two default bench invocations
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
SECONDS=0; timeout 900 python bench.py > gpurun_out/r1j_bench_default.json 2> gpurun_out/r1j_bench_default.err; echo "bench rc=$?"; tail -2 gpurun_out/r1j_bench_default.err; python tools/bench_summary.py < gpurun_out/r1j_bench_default.json
echo "default bench wall $SECONDS s"; SECONDS=0; timeout 900 python bench.py --impl reference > gpurun_out/r1j_bench_reference.json 2> gpurun_out/r1j_bench_reference.err; echo "ref rc=$?"; tail -1 gpurun_out/r1j_bench_reference.err; cut -c1-300 gpurun_out/r1j_bench_reference.json
echo "reference arm wall $SECONDS s"; timeout 900 python bench.py --mode realign --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/r1j_bench_realign.json 2> gpurun_out/r1j_bench_realign.err; echo "realign rc=$?"; python tools/bench_summary.py < gpurun_out/r1j_bench_realign.json
