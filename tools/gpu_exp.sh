# quick experiment matrix for the 5-state wavefront kernels (coefficient source x warps per block)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "golden or batched or wide" > gpurun_out/pytest_exp.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_exp.log
for cm in mixed smem const; do for nw in 4 8 16; do
  echo "== MAB_COEF=$cm MAB_W5_WARPS=$nw"
  MAB_COEF=$cm MAB_W5_WARPS=$nw timeout 300 python bench.py --steps 3 --warmup 2 --reads 200000 --no-e2e --no-cpu-baseline 2>&1 | python tools/bench_summary.py
done; done
