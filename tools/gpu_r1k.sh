# EM bench lines after the table-driven exp in k_expect + ncu of the kernel
mkdir -p gpurun_out
run() { name=$1; shift; timeout 900 python bench.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"; tail -2 gpurun_out/$name.err; python tools/bench_summary.py < gpurun_out/$name.json; }
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
run r1k_bench_em --steps 3 --warmup 3 --mode em --reads 300000 --no-cpu-baseline
run r1k_bench_trna3_em --steps 3 --warmup 3 --mode em --states 3 --reads 300000 --no-cpu-baseline
run r1k_bench_long_em --steps 3 --warmup 3 --mode em --workload long --no-cpu-baseline
CMD="python bench.py --steps 1 --warmup 1 --reads 200000 --mode em --no-e2e --no-cpu-baseline"
ncu --set full --clock-control none --import-source on -k regex:'k_expect' -s 2 -c 1 -o gpurun_out/r1k_expect -f $CMD > gpurun_out/ncu_x.log 2>&1; echo "ncu expect rc=$?"
