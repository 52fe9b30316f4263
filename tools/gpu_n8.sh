# N = 8 check of the torchrun path the driver uses for the scaling run
mkdir -p gpurun_out
nvidia-smi --query-gpu=name --format=csv,noheader | wc -l; nproc; free -g | head -2
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 8 --steps 3 --warmup 3 > gpurun_out/r1i_bench_trna_n8.json 2> gpurun_out/n8.err; echo "n8 rc=$?"; tail -3 gpurun_out/n8.err; python tools/bench_summary.py < gpurun_out/r1i_bench_trna_n8.json
