#!/bin/bash
mkdir -p gpurun_out
N=$1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N > gpurun_out/r2an_bench_n$N.json 2> gpurun_out/r2an_bench_n$N.err; echo "bench n$N rc=$?"; tail -3 gpurun_out/r2an_bench_n$N.err
python tools/bench_summary.py < gpurun_out/r2an_bench_n$N.json
python - <<PY
import json
j=json.loads(open("gpurun_out/r2an_bench_n$N.json").read().strip().splitlines()[-1])
print("e2e", j["e2e"]["value"], "rl", j["e2e_realign"] and j["e2e_realign"]["value"], "em", j["em"], "strong", j["strong"])
PY
