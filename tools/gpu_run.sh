#!/bin/bash
# One parametrised GPU session (replaces the round-1 tools/gpu_*.sh one-offs).  Usage under gpurun:
#   tools/gpu_run.sh <tag> [tests] [smoke] [bench:<name>:<bench.py args>] ... [ncu_list:<args>] [ncu_full:<regex>:<args>]
# Everything lands in gpurun_out/<tag>_*.
tag=$1; shift
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > gpurun_out/${tag}_gpu.txt; nproc >> gpurun_out/${tag}_gpu.txt
for step in "$@"; do
  kind=${step%%:*}; rest=${step#*:}
  case $kind in
    tests) eval "timeout 2400 python -m pytest tests -m gpu -x -q ${rest#tests}" > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/${tag}_pytest.log;;  # eval: -k "a or b" keeps its quotes
    smoke) timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${tag}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/${tag}_smoke.log;;
    bench) name=${rest%%:*}; args=${rest#*:}
      timeout 1200 python bench.py $args > gpurun_out/${tag}_bench_$name.json 2> gpurun_out/${tag}_bench_$name.err; echo "bench $name rc=$?"; tail -2 gpurun_out/${tag}_bench_$name.err
      python tools/bench_summary.py < gpurun_out/${tag}_bench_$name.json;;
    ncu_list) timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${tag}_launches.csv python bench.py $rest > gpurun_out/${tag}_ncu_list.log 2>&1; echo "ncu list rc=$?";;
    ncu_full) regex=${rest%%:*}; args=${rest#*:}
      timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"$regex" -s 2 -c 2 -o gpurun_out/${tag}_full -f python bench.py $args > gpurun_out/${tag}_ncu_full.log 2>&1; echo "ncu full rc=$?"
      ncu -i gpurun_out/${tag}_full.ncu-rep --page raw --csv > gpurun_out/${tag}_full_raw.csv 2>/dev/null;;
    sh) bash -c "$rest";;
  esac
done
