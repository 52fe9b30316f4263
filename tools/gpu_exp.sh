mkdir -p gpurun_out
run() { name=$1; shift; timeout 900 python bench.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"; tail -2 gpurun_out/$name.err; python tools/bench_summary.py < gpurun_out/$name.json; }
run xt_em_nb6 --steps 3 --warmup 3 --mode em --reads 300000 --no-cpu-baseline --no-e2e
run xt_em3_nb6 --steps 3 --warmup 3 --mode em --states 3 --reads 300000 --no-cpu-baseline --no-e2e
run xt_emlong_nb6 --steps 3 --warmup 3 --mode em --workload long --no-cpu-baseline --no-e2e
