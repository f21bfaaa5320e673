mkdir -p gpurun_out; : > gpurun_out/mqexp.log
for nl in 0 64; do
  echo "== J2KCUDA_MQ_NLONG=$nl" >> gpurun_out/mqexp.log
  J2KCUDA_MQ_NLONG=$nl python bench.py --steps 3 --warmup 2 --no-cpu-baseline 2>/dev/null | python tools/bench_summary.py /dev/stdin 2>/dev/null | head -4 | cut -c1-300 >> gpurun_out/mqexp.log
done
cat gpurun_out/mqexp.log
