# full check after Tier-1 changes: all GPU parity tests, stage times of both directions, bench
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/c_tests.log
cat gpurun_out/c_tests.log
python tools/stage_times.py 2>&1 | head -12 > gpurun_out/c_stage_enc.log; cat gpurun_out/c_stage_enc.log
python tools/stage_times.py decode > gpurun_out/c_stage_dec.log 2>&1; cat gpurun_out/c_stage_dec.log
python bench.py --no-cpu-baseline > gpurun_out/c_bench.json 2> gpurun_out/c_bench.err; tail -3 gpurun_out/c_bench.err
