# full GPU parity + C3 timing + per-config throughput
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > gpurun_out/h_tests.log; cat gpurun_out/h_tests.log
python tools/stage_times.py c3 2>&1 | grep -E "host_t2|DECODE|host_parse" > gpurun_out/h_stage_c3.log; cat gpurun_out/h_stage_c3.log
python tools/bench_configs.py --reps 2 --workers 16 > gpurun_out/h_configs_1gpu.jsonl 2> gpurun_out/h_configs_1gpu.err
tail -3 gpurun_out/h_configs_1gpu.err; cut -c1-330 gpurun_out/h_configs_1gpu.jsonl
