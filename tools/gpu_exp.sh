mkdir -p gpurun_out
MAB_LA_MODE=1 timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/pytest_m1.log 2>&1; echo "pytest mode1 rc=$?"; tail -3 gpurun_out/pytest_m1.log
run() { name=$1; shift; timeout 900 python bench.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"; tail -2 gpurun_out/$name.err; python tools/bench_summary.py < gpurun_out/$name.json; }
MAB_LA_MODE=0 run trna_m0 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e
MAB_LA_MODE=1 run trna_m1 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e
MAB_LA_MODE=0 run trna3_m0 --steps 3 --warmup 3 --states 3 --no-cpu-baseline --no-e2e
MAB_LA_MODE=1 run trna3_m1 --steps 3 --warmup 3 --states 3 --no-cpu-baseline --no-e2e
