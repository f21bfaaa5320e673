# 2-GPU runs: bench.py and the per-configuration throughput
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --no-cpu-baseline > gpurun_out/i_bench_2gpu.json 2> gpurun_out/i_bench_2gpu.err
tail -2 gpurun_out/i_bench_2gpu.err; cut -c1-300 gpurun_out/i_bench_2gpu.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 tools/bench_configs.py --reps 2 --workers 16 > gpurun_out/i_configs_2gpu.jsonl 2> gpurun_out/i_configs_2gpu.err
tail -3 gpurun_out/i_configs_2gpu.err; cut -c1-260 gpurun_out/i_configs_2gpu.jsonl
