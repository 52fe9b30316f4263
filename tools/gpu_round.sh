mkdir -p gpurun_out
nvidia-smi > gpurun_out/smi.txt 2>&1
nproc > gpurun_out/nproc.txt
./tools/pipe_peaks > gpurun_out/pipe_peaks.json 2> gpurun_out/pipe_peaks.err; echo "peaks rc=$?"
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_trna.json 2> gpurun_out/bench_trna.err; echo "bench rc=$?"; cat gpurun_out/bench_trna.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cat gpurun_out/bench_ref.json
timeout 600 python bench.py --mode em --steps 3 --warmup 3 --reads 300000 --no-cpu-baseline > gpurun_out/bench_em.json 2> gpurun_out/bench_em.err; echo "em rc=$?"; cat gpurun_out/bench_em.json
timeout 600 python bench.py --workload long --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_long.json 2> gpurun_out/bench_long.err; echo "long rc=$?"; cat gpurun_out/bench_long.json
PROF="python bench.py --steps 1 --warmup 1 --reads 200000 --no-e2e --no-cpu-baseline"
$PROF > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 100 --csv --log-file gpurun_out/launches.csv $PROF > gpurun_out/ncu_launch.log 2>&1; echo "ncu1 rc=$?"
$PROF > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'k_(forward|backward)5' -s 2 -c 2 -o gpurun_out/prof_r1 $PROF > gpurun_out/ncu_full.log 2>&1; echo "ncu2 rc=$?"
ls -la gpurun_out
