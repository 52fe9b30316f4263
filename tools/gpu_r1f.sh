# r1f: state capture after the wide-band / realign work: bench lines, ncu launch lists, ncu --set full of the DP kernels
mkdir -p gpurun_out
run() { name=$1; shift; timeout 900 python bench.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"; tail -2 gpurun_out/$name.err; python tools/bench_summary.py < gpurun_out/$name.json; }
run r1f_bench_trna --steps 5 --warmup 3
run r1f_bench_em --steps 3 --warmup 3 --mode em --reads 300000 --no-cpu-baseline
run r1f_bench_long --steps 3 --warmup 3 --workload long --no-cpu-baseline
run r1f_bench_long_em --steps 3 --warmup 3 --workload long --mode em --no-cpu-baseline
run r1f_bench_realign --steps 3 --warmup 3 --mode realign --no-cpu-baseline
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r1f_bench_ref.json 2> gpurun_out/r1f_bench_ref.err; echo "ref rc=$?"
CMD="python bench.py --steps 1 --warmup 1 --reads 200000 --mode realign --no-e2e --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1f_launches_trna_realign.csv $CMD > gpurun_out/ncu_l.log 2>&1; echo "ncu list rc=$?"
CMDL="python bench.py --steps 1 --warmup 1 --workload long --reads 1000 --no-e2e --no-cpu-baseline"
$CMDL > gpurun_out/plainl.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1f_launches_long.csv $CMDL > gpurun_out/ncu_ll.log 2>&1; echo "ncu list long rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'k_(forward|backward)5' -s 2 -c 2 -o gpurun_out/r1f_dp_long -f $CMDL > gpurun_out/ncu_fl.log 2>&1; echo "ncu full long rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'k_(posterior_diag|rl_chain|rl_weights|totals)' -s 4 -c 4 -o gpurun_out/r1f_post_trna -f $CMD > gpurun_out/ncu_p.log 2>&1; echo "ncu post rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'k_(forward|backward)5' -s 2 -c 2 -o gpurun_out/r1f_dp_trna -f $CMD > gpurun_out/ncu_f.log 2>&1; echo "ncu full trna rc=$?"
ls -la gpurun_out | head -50
