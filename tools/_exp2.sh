mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_shard_nccl.py tests/test_realign_sam.py -m gpu -x -q > gpurun_out/r2k_pytest_n2.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2k_pytest_n2.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/r2k_bench_n2.json 2> gpurun_out/r2k_bench_n2.err; echo "bench rc=$?"; tail -3 gpurun_out/r2k_bench_n2.err; python tools/bench_summary.py < gpurun_out/r2k_bench_n2.json
