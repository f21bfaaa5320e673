# per-configuration throughput (C1-C5), 1 GPU
mkdir -p gpurun_out
python tools/bench_configs.py --reps 2 > gpurun_out/e_configs_1gpu.jsonl 2> gpurun_out/e_configs_1gpu.err
tail -3 gpurun_out/e_configs_1gpu.err; cat gpurun_out/e_configs_1gpu.jsonl
