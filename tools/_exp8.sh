mkdir -p gpurun_out
N=$1
run() { name=$1; shift; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@" > gpurun_out/r2w_bench_n${N}_$name.json 2> gpurun_out/r2w_bench_n${N}_$name.err; echo "$name rc=$?"; python tools/bench_summary.py < gpurun_out/r2w_bench_n${N}_$name.json; }
run c32 --steps 3 --warmup 3 --no-em-leg --no-strong-leg --no-realign-leg --int32-pairs
run c16 --steps 3 --warmup 3
