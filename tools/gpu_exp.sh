mkdir -p gpurun_out
run() { name=$1; shift; timeout 900 python bench.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"; tail -2 gpurun_out/$name.err; python tools/bench_summary.py < gpurun_out/$name.json; python -c "
import json,sys
d=json.loads(open('gpurun_out/$name.json').read().strip().splitlines()[-1]); e=d.get('e2e') or {}
print('   e2e ms', e.get('ms_per_step'), {k: round(v,3) for k,v in (e.get('host_wall_s') or {}).items()})"; }
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
run ch_post --steps 5 --warmup 3 --no-cpu-baseline
run ch_post2 --steps 5 --warmup 3 --no-cpu-baseline
run ch_realign --steps 4 --warmup 3 --no-cpu-baseline --mode realign
run ch_long --steps 3 --warmup 3 --no-cpu-baseline --workload long
