#!/bin/bash
mkdir -p gpurun_out
A="--steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-em-leg --no-realign-leg --no-strong-leg"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_bands|k_schedule|k_regions" -s 8 -c 4 -o gpurun_out/r2ac_plan -f python bench.py $A > gpurun_out/r2ac_ncu.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/r2ac_plan.ncu-rep --page raw --csv > gpurun_out/r2ac_plan_raw.csv 2>/dev/null
python tools/ncu_pick.py gpurun_out/r2ac_plan_raw.csv
timeout 900 ncu --set full --clock-control none -k regex:"k_bands|k_schedule|k_regions" -s 4 -c 4 -o gpurun_out/r2ac_plan_long -f python bench.py $A --workload long --expansion 100 --reads 1200 > gpurun_out/r2ac_ncu_long.log 2>&1; echo "ncu long rc=$?"
ncu -i gpurun_out/r2ac_plan_long.ncu-rep --page raw --csv > gpurun_out/r2ac_plan_long_raw.csv 2>/dev/null
python tools/ncu_pick.py gpurun_out/r2ac_plan_long_raw.csv
