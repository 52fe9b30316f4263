mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu.log
run() { name=$1; shift; timeout 900 python bench.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"; tail -2 gpurun_out/$name.err; python tools/bench_summary.py < gpurun_out/$name.json; python - <<PY
import json
d=json.loads(open("gpurun_out/$name.json").read().strip().splitlines()[-1]); e=d["e2e"]; print(round(e["value"],3), round(e["ms_per_step"],1), e["host_wall_s"])
PY
}
MAB_POSTERIOR_SYNC=1 run e2e_sync --steps 3 --warmup 3 --e2e-steps 3 --no-cpu-baseline
run e2e_async --steps 3 --warmup 3 --e2e-steps 3 --no-cpu-baseline
run e2e_realign --steps 3 --warmup 3 --e2e-steps 3 --mode realign --no-cpu-baseline
run e2e_em --steps 3 --warmup 3 --e2e-steps 3 --mode em --reads 300000 --no-cpu-baseline
run long_realign --steps 2 --warmup 3 --workload long --mode realign --no-cpu-baseline
