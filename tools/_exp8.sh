mkdir -p gpurun_out
N=$1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/r2l_bench_n$N.json 2> gpurun_out/r2l_bench_n$N.err; echo "bench rc=$?"; tail -3 gpurun_out/r2l_bench_n$N.err | cut -c1-300; python tools/bench_summary.py < gpurun_out/r2l_bench_n$N.json; nproc; df -h /dev/shm | tail -1
