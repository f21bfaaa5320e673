mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 tools/bench_configs.py --reps 3 --workers 16 --only C4 --frames8k 6 > gpurun_out/k_c4_2gpu.jsonl 2> gpurun_out/k_c4_2gpu.err
tail -3 gpurun_out/k_c4_2gpu.err; cat gpurun_out/k_c4_2gpu.jsonl
python tools/bench_configs.py --reps 3 --only C4 --frames8k 6 > gpurun_out/k_c4_1gpu.jsonl 2>/dev/null; cat gpurun_out/k_c4_1gpu.jsonl
