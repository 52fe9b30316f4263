# N = 4 check of the torchrun path the driver uses for the scaling run (posterior and EM, the latter with the NCCL all-reduce)
mkdir -p gpurun_out
nvidia-smi --query-gpu=name --format=csv,noheader | head -8; nproc
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 4 --steps 3 --warmup 3 > gpurun_out/r1i_bench_trna_n4.json 2> gpurun_out/n4.err; echo "n4 rc=$?"; tail -3 gpurun_out/n4.err; python tools/bench_summary.py < gpurun_out/r1i_bench_trna_n4.json
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus 4 --steps 3 --warmup 3 --mode em --reads 300000 > gpurun_out/r1i_bench_em_n4.json 2> gpurun_out/n4em.err; echo "n4 em rc=$?"; tail -3 gpurun_out/n4em.err; python tools/bench_summary.py < gpurun_out/r1i_bench_em_n4.json
