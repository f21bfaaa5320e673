mkdir -p gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -6 > gpurun_out/r2g_gpu_tests.log
python bench.py > gpurun_out/r2g_bench.json 2> gpurun_out/r2g_bench.err
cat gpurun_out/r2g_gpu_tests.log; python tools/bench_summary.py gpurun_out/r2g_bench.json
