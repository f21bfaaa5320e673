#!/bin/bash
# Multi-GPU evidence (run through `gpurun --gpus N`): tools/gpu_multi.sh N [tag].  bench.py over N ranks, the multi-device tests on
# real devices, configurations C2 / C4 / C5 over N ranks and the single-process multi-device arms of the C ABI.
N=${1:-2}; tag=${2:-r2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
nvidia-smi -L > gpurun_out/${tag}_${N}gpu_devices.log; nproc >> gpurun_out/${tag}_${N}gpu_devices.log
timeout 600 python -m pytest tests/test_gpu_sequence.py -m gpu -q 2>&1 | tail -5 > gpurun_out/${tag}_${N}gpu_tests.log
timeout 800 $TR bench.py --gpus $N --no-cpu-baseline > gpurun_out/${tag}_bench_${N}gpu.json 2> gpurun_out/${tag}_bench_${N}gpu.err
timeout 800 $TR tools/bench_configs.py --only C2,C4,C5 --reps 2 > gpurun_out/${tag}_configs_${N}gpu.jsonl 2> gpurun_out/${tag}_configs_${N}gpu.err
timeout 600 python tools/bench_configs.py --only C4api,C5api --reps 2 --devices $N >> gpurun_out/${tag}_configs_${N}gpu.jsonl 2>> gpurun_out/${tag}_configs_${N}gpu.err
cat gpurun_out/${tag}_${N}gpu_tests.log
python tools/bench_summary.py gpurun_out/${tag}_bench_${N}gpu.json
cut -c1-400 gpurun_out/${tag}_configs_${N}gpu.jsonl
tail -3 gpurun_out/${tag}_configs_${N}gpu.err
