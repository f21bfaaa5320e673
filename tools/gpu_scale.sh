#!/bin/bash
# Scaling evidence on one N-GPU box (run through `gpurun --gpus N`, charged N x the box time, so kept short):
#   bench.py over N ranks; C4 (tiles of an 8K frame) through j2kcuda_tiles_* from ONE process on 1, 2, 4 .. N devices;
#   C5 through j2kcuda_sequence_create_multi on N devices.
N=${1:-8}; tag=${2:-r2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
(nvidia-smi -L; nproc) > gpurun_out/${tag}_${N}gpu_devices.log
timeout 600 $TR bench.py --gpus $N --no-cpu-baseline > gpurun_out/${tag}_bench_${N}gpu.json 2> gpurun_out/${tag}_bench_${N}gpu.err
out=gpurun_out/${tag}_scale_${N}gpu.jsonl; : > $out
for d in 1 2 4 8; do
  [ $d -le $N ] && timeout 300 python tools/bench_configs.py --only C4api --reps 2 --frames8k 3 --devices $d >> $out 2>> gpurun_out/${tag}_scale_${N}gpu.err
done
timeout 300 python tools/bench_configs.py --only C5api --reps 2 --devices $N >> $out 2>> gpurun_out/${tag}_scale_${N}gpu.err
python tools/bench_summary.py gpurun_out/${tag}_bench_${N}gpu.json 2>/dev/null | head -6
cut -c1-330 $out
tail -3 gpurun_out/${tag}_scale_${N}gpu.err
