mkdir -p gpurun_out
run() { name=$1; shift; timeout 900 python bench.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"; tail -2 gpurun_out/$name.err; python tools/bench_summary.py < gpurun_out/$name.json; python -c "
import json,sys
d=json.loads(open('gpurun_out/$name.json').read().strip().splitlines()[-1]); e=d.get('e2e') or {}
print('   e2e ms', e.get('ms_per_step'), {k: round(v,3) for k,v in (e.get('host_wall_s') or {}).items()}, d['config'].get('chunks'))"; }
MAB_CHUNK_RAMP=4 run ramp4 --steps 5 --warmup 3 --no-cpu-baseline
run ramp842 --steps 5 --warmup 3 --no-cpu-baseline
MAB_CHUNK_RAMP=16,8,4,2 run ramp16 --steps 5 --warmup 3 --no-cpu-baseline
MAB_CHUNK_RAMP=4 run ramp4_rl --steps 4 --warmup 3 --no-cpu-baseline --mode realign
run ramp842_rl --steps 4 --warmup 3 --no-cpu-baseline --mode realign
run ramp842_em --steps 4 --warmup 3 --no-cpu-baseline --mode em --reads 300000
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
