mkdir -p gpurun_out
cp marginalign-trna_b200/libmargin_b200.so /tmp/lib6.so
for v in 6 4 5 7; do
  if [ $v = 6 ]; then cp /tmp/lib6.so marginalign-trna_b200/libmargin_b200.so; else cp tools/_variants/libmargin_minb$v.so marginalign-trna_b200/libmargin_b200.so; fi
  timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-em-leg --no-strong-leg --no-realign-leg > gpurun_out/r2o_bench_minb$v.json 2>gpurun_out/r2o_bench_minb$v.err; echo "MINB=$v"; python tools/bench_summary.py < gpurun_out/r2o_bench_minb$v.json
  timeout 600 python bench.py --steps 3 --warmup 3 --mode em --reads 300000 --no-cpu-baseline --no-e2e --no-em-leg --no-strong-leg --no-realign-leg > gpurun_out/r2o_bench_em_minb$v.json 2>gpurun_out/r2o_bench_em_minb$v.err; python tools/bench_summary.py < gpurun_out/r2o_bench_em_minb$v.json
done
cp /tmp/lib6.so marginalign-trna_b200/libmargin_b200.so
