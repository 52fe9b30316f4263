mkdir -p gpurun_out
./tools/smem_probe > gpurun_out/smem_probe.json; cat gpurun_out/smem_probe.json
timeout 900 python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/bench_trna.json 2> gpurun_out/bench_trna.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_trna.err; python tools/bench_summary.py < gpurun_out/bench_trna.json
python -c "
import json; d=json.load(open('gpurun_out/bench_trna.json')); print(d['e2e'])"
