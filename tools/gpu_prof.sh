mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "device_planner or host_and_device" > gpurun_out/pytest_plan.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_plan.log
PROF="python bench.py --steps 1 --warmup 1 --reads 200000 --no-e2e --no-cpu-baseline"
$PROF > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'k_(forward5|backward5|posterior_count|emit)' -s 4 -c 4 -o gpurun_out/prof_r1b $PROF > gpurun_out/ncu_full.log 2>&1; echo "ncu rc=$?"
