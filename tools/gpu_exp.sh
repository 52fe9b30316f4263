mkdir -p gpurun_out
run() { name=$1; shift; timeout 900 python bench.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"; tail -2 gpurun_out/$name.err; python tools/bench_summary.py < gpurun_out/$name.json; python -c "
import json,sys
d=json.loads(open('gpurun_out/$name.json').read().strip().splitlines()[-1]); e=d.get('e2e') or {}
print('   e2e ms', e.get('ms_per_step'), {k: round(v,3) for k,v in (e.get('host_wall_s') or {}).items()})"; }
timeout 600 python -m pytest tests/test_gpu_realign.py -m gpu -x -q > gpurun_out/pytest_rl.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_rl.log
run rl_lazy --steps 4 --warmup 3 --no-cpu-baseline --mode realign
run rl_lazy2 --steps 4 --warmup 3 --no-cpu-baseline --mode realign
CMD="python bench.py --steps 1 --warmup 1 --reads 200000 --mode em --no-e2e --no-cpu-baseline"
ncu --set full --clock-control none --import-source on -k regex:'k_expect' -s 2 -c 1 -o gpurun_out/r1i_expect -f $CMD > gpurun_out/ncu_x.log 2>&1; echo "ncu expect rc=$?"
